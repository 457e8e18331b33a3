// bvh_build.cu — BVH construction on the device.
//
// Replaces chroma/bvh/grid.py:11-95 (make_recursive_grid_bvh: host argsort + numpy grouping
// loop with a PCIe round trip per layer) and the kernels it drives in chroma/cuda/bvh.cu:
// make_leaves :148-203, make_parents_detailed :269-308, copy_and_offset :364-384,
// collapse_child :530-543, plus chroma/gpu/bvh.py:17-130,239-267.  Everything stays in HBM:
// Morton codes are radix-sorted on device (stable, so ties keep triangle order), the
// "shift until mean group size >= target_degree" search is one pass (a histogram of the
// highest differing bit between neighbouring codes gives the group count for every shift at
// once), groups are found with a scan and split at 15 children exactly as grid.py:48-71.
// Divisions use IEEE rounding (the reference kernels are built with --use_fast_math), so
// the result is bit-identical to the CPU restatement in oracle/; see DESIGN.md.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include "common.cuh"

#define MAX_CHILD 15u
#define CHILD_BITS 28

// order-preserving float <-> uint mapping for atomicMin/Max
__device__ __forceinline__ uint32_t f2ord(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
static inline float ord2f(uint32_t u) {
    uint32_t v = (u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u;
    float f;
    memcpy(&f, &v, 4);
    return f;
}

__global__ void bounds_kernel(const float *v, uint64_t nv, uint32_t *out /* min xyz, max xyz */) {
    float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += stride) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            float x = v[3 * i + k];
            mn[k] = fminf(mn[k], x);
            mx[k] = fmaxf(mx[k], x);
        }
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
        for (int o = 16; o > 0; o >>= 1) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(out + k, f2ord(mn[k]));
            atomicMax(out + 3 + k, f2ord(mx[k]));
        }
    }
}

// bvh.cu:42-52
__device__ __forceinline__ unsigned long long spread3_16(unsigned int input) {
    unsigned long long x = input;
    x = (x | (x << 16)) & 0x00000000FF0000FFull;
    x = (x | (x << 8)) & 0x000000F00F00F00Full;
    x = (x | (x << 4)) & 0x00000C30C30C30C3ull;
    x = (x | (x << 2)) & 0x0000249249249249ull;
    return x;
}
// bvh.cu:65-76 (truncating quantisation).  Plain operators on purpose: this library is built
// with --use_fast_math like the reference (chroma/gpu/tools.py:12), so `/` is the same
// div.approx the reference kernel executes and the quotients truncate identically
// (tests/test_reference_gpu.py::test_make_leaves_bit_exact compares against the reference cubin).
__device__ __forceinline__ unsigned quantize(float v, float origin, float scale) {
    return (unsigned)((v - origin) / scale);
}

// bvh.cu:148-203 make_leaves
__global__ void make_leaves_kernel(uint64_t nt, const uint32_t *tri, const float *vert, float3 wo,
                                   float ws, uint4 *leaf, unsigned long long *morton) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < nt; t += stride) {
        uint32_t i0 = tri[3 * t], i1 = tri[3 * t + 1], i2 = tri[3 * t + 2];
        float3 a = make_float3(vert[3 * (size_t)i0], vert[3 * (size_t)i0 + 1], vert[3 * (size_t)i0 + 2]);
        float3 b = make_float3(vert[3 * (size_t)i1], vert[3 * (size_t)i1 + 1], vert[3 * (size_t)i1 + 2]);
        float3 c = make_float3(vert[3 * (size_t)i2], vert[3 * (size_t)i2 + 1], vert[3 * (size_t)i2 + 2]);
        float3 lo = make_float3(fminf(fminf(a.x, b.x), c.x), fminf(fminf(a.y, b.y), c.y), fminf(fminf(a.z, b.z), c.z));
        float3 hi = make_float3(fmaxf(fmaxf(a.x, b.x), c.x), fmaxf(fmaxf(a.y, b.y), c.y), fmaxf(fmaxf(a.z, b.z), c.z));
        float3 ce = make_float3(((a.x + b.x) + c.x) / 3.0f, ((a.y + b.y) + c.y) / 3.0f,
                                ((a.z + b.z) + c.z) / 3.0f);
        unsigned qlx = quantize(lo.x, wo.x, ws), qly = quantize(lo.y, wo.y, ws), qlz = quantize(lo.z, wo.z, ws);
        if (qlx > 0) qlx--;
        if (qly > 0) qly--;
        if (qlz > 0) qlz--;
        unsigned qux = quantize(hi.x, wo.x, ws) + 1, quy = quantize(hi.y, wo.y, ws) + 1, quz = quantize(hi.z, wo.z, ws) + 1;
        unsigned qcx = quantize(ce.x, wo.x, ws), qcy = quantize(ce.y, wo.y, ws), qcz = quantize(ce.z, wo.z, ws);
        unsigned long long m = spread3_16(qcx) | (spread3_16(qcy) << 1) | (spread3_16(qcz) << 2);
        leaf[t] = make_uint4(qlx | (qux << 16), qly | (quy << 16), qlz | (quz << 16), (uint32_t)t);
        morton[t] = m;
    }
}

__global__ void iota_kernel(uint32_t *a, uint64_t n) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) a[i] = (uint32_t)i;
}
__global__ void gather_nodes_kernel(const uint4 *in, const uint32_t *idx, uint4 *out, uint64_t n) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = in[idx[i]];
}

// histogram of the highest differing bit between neighbouring (sorted) codes
__global__ void diffbit_hist_kernel(const unsigned long long *codes, uint64_t n, unsigned long long *hist) {
    __shared__ unsigned int sh[64];
    if (threadIdx.x < 64) sh[threadIdx.x] = 0;
    __syncthreads();
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x + 1; i < n; i += stride) {
        unsigned long long x = codes[i] ^ codes[i - 1];
        if (x) atomicAdd(&sh[63 - __clzll((long long)x)], 1u);
    }
    __syncthreads();
    if (threadIdx.x < 64 && sh[threadIdx.x]) atomicAdd(hist + threadIdx.x, (unsigned long long)sh[threadIdx.x]);
}

// shift codes and flag group heads
__global__ void shift_heads_kernel(unsigned long long *codes, uint64_t n, int shift, uint32_t *head) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        unsigned long long c = codes[i] >> shift;
        unsigned long long p = i ? (codes[i - 1] >> shift) : 0ull;
        head[i] = (i == 0 || c != p) ? 1u : 0u;
    }
}
__global__ void apply_shift_kernel(unsigned long long *codes, uint64_t n, int shift) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) codes[i] >>= shift;
}
// group_start[gid] = i for every head i (gid = exclusive scan of head)
__global__ void group_start_kernel(const uint32_t *head, const uint32_t *gid, uint64_t n, uint32_t *group_start) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        if (head[i]) group_start[gid[i]] = (uint32_t)i;
}
// parents per group: max(1, ceil(size/15))  (grid.py:48-71)
__global__ void group_parents_kernel(const uint32_t *group_start, uint64_t ngroups, uint64_t n, uint32_t *nparents) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += stride) {
        uint32_t s = group_start[g];
        uint32_t e = (g + 1 < ngroups) ? group_start[g + 1] : (uint32_t)n;
        uint32_t size = e - s;
        nparents[g] = (size + MAX_CHILD - 1) / MAX_CHILD;
    }
}
// bvh.cu:269-308 make_parents_detailed, fused with the first_child/nchild bookkeeping
__global__ void make_parents_kernel(const uint32_t *group_start, const uint32_t *parent_off, uint64_t ngroups,
                                    uint64_t n, const uint4 *child_nodes, const unsigned long long *codes,
                                    uint4 *parent_nodes, unsigned long long *parent_codes) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += stride) {
        uint32_t s = group_start[g];
        uint32_t e = (g + 1 < ngroups) ? group_start[g + 1] : (uint32_t)n;
        uint32_t po = parent_off[g];
        unsigned long long code = codes[s];
        for (uint32_t first = s, k = 0; first < e; first += MAX_CHILD, k++) {
            uint32_t nchild = min(MAX_CHILD, e - first);
            uint4 c0 = child_nodes[first];
            uint3 lo = make_uint3(c0.x & 0xFFFF, c0.y & 0xFFFF, c0.z & 0xFFFF);
            uint3 hi = make_uint3(c0.x >> 16, c0.y >> 16, c0.z >> 16);
            for (uint32_t i = 1; i < nchild; i++) {
                uint4 c = child_nodes[first + i];
                lo.x = min(lo.x, c.x & 0xFFFF); lo.y = min(lo.y, c.y & 0xFFFF); lo.z = min(lo.z, c.z & 0xFFFF);
                hi.x = max(hi.x, c.x >> 16); hi.y = max(hi.y, c.y >> 16); hi.z = max(hi.z, c.z >> 16);
            }
            parent_nodes[po + k] = make_uint4(hi.x << 16 | lo.x, hi.y << 16 | lo.y, hi.z << 16 | lo.z,
                                              (nchild << CHILD_BITS) | first);
            parent_codes[po + k] = code;
        }
    }
}
// bvh.cu:364-384 copy_and_offset
__global__ void copy_and_offset_kernel(const uint4 *src, uint4 *dst, uint64_t n, uint32_t child_offset) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        uint4 v = src[i];
        uint32_t nchild = v.w >> CHILD_BITS, child = v.w & 0x0FFFFFFFu;
        v.w = (nchild << CHILD_BITS) | (child + child_offset);
        dst[i] = v;
    }
}
// bvh.cu:530-543 collapse_child
__global__ void collapse_child_kernel(uint64_t start, uint64_t end, uint4 *node) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = start + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < end; i += stride) {
        uint4 v = node[i];
        if ((v.w >> CHILD_BITS) == 1) node[i] = node[v.w & 0x0FFFFFFFu];
    }
}
// standalone make_parents_detailed for the chroma/gpu/bvh.py helper API
__global__ void make_parents_detailed_kernel(uint64_t nparent, const uint4 *child_nodes, uint4 *parent_nodes,
                                             const int *first_children, const int *nchildren) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < nparent; p += stride) {
        uint32_t first = first_children[p], nchild = nchildren[p];
        uint4 c0 = child_nodes[first];
        uint3 lo = make_uint3(c0.x & 0xFFFF, c0.y & 0xFFFF, c0.z & 0xFFFF);
        uint3 hi = make_uint3(c0.x >> 16, c0.y >> 16, c0.z >> 16);
        for (uint32_t i = 1; i < nchild; i++) {
            uint4 c = child_nodes[first + i];
            lo.x = min(lo.x, c.x & 0xFFFF); lo.y = min(lo.y, c.y & 0xFFFF); lo.z = min(lo.z, c.z & 0xFFFF);
            hi.x = max(hi.x, c.x >> 16); hi.y = max(hi.y, c.y >> 16); hi.z = max(hi.z, c.z >> 16);
        }
        parent_nodes[p] = make_uint4(hi.x << 16 | lo.x, hi.y << 16 | lo.y, hi.z << 16 | lo.z,
                                     (nchild << CHILD_BITS) | first);
    }
}

#define CK(expr)                                                                           \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess) {                                                           \
            chroma_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            rc = CHROMA_ERR_CUDA;                                                          \
            goto done;                                                                     \
        }                                                                                  \
    } while (0)

struct Layer { uint64_t off, count; };  // position in the `upper` build buffer

// Either (vertices, triangles) or ready-made leaves + 48-bit Morton codes + world coordinates
// (chroma_bvh_build_from_leaves: the host part of grid.py on given leaves, what the parity tests
// compare with the reference's grid.py run on the same leaves).
static int bvh_build_impl(chroma_ctx *c, const float *vertices, uint64_t nv,
                          const uint32_t *triangles, uint64_t nt, const uint32_t *leaves_in,
                          const uint64_t *codes_in, const float *origin_in, float scale_in,
                          int target_degree, chroma_bvh **out) {
    CTX_BIND(c);
    ARG_CHECK(out && nt > 0 && nt < (1ull << 28));
    ARG_CHECK((vertices && triangles && nv > 0) || (leaves_in && codes_in && origin_in));
    ARG_CHECK(target_degree >= 1);
    int rc = 0;
    cudaStream_t st = c->stream;
    float *d_vert = nullptr; uint32_t *d_tri = nullptr;
    uint4 *d_leaf = nullptr, *d_leaf_sorted = nullptr, *d_upper = nullptr;
    unsigned long long *d_code_a = nullptr, *d_code_b = nullptr, *d_hist = nullptr;
    uint32_t *d_idx_a = nullptr, *d_idx_b = nullptr, *d_head = nullptr, *d_gid = nullptr,
             *d_gstart = nullptr, *d_npar = nullptr, *d_poff = nullptr, *d_bounds = nullptr;
    void *d_temp = nullptr; size_t temp_cap = 0;
    chroma_bvh *bvh = new chroma_bvh();
    bvh->ctx = c;
    std::vector<Layer> layers;  // bottom-up, excluding the leaf layer
    const uint64_t upper_cap = nt + 1024;
    uint64_t upper_used = 0;
    unsigned grid = grid_for(nt, 256, c, 8);
    unsigned long long h_hist[64];
    uint32_t h_bounds[6];
    if (leaves_in) {
        CK(cudaMalloc(&d_hist, 64 * 8));
        for (int k = 0; k < 3; k++) bvh->world_origin[k] = origin_in[k];
        bvh->world_scale = scale_in;
        CK(cudaMalloc(&d_leaf, nt * 16)); CK(cudaMalloc(&d_leaf_sorted, nt * 16));
        CK(cudaMalloc(&d_code_a, nt * 8)); CK(cudaMalloc(&d_code_b, nt * 8));
        CK(cudaMalloc(&d_idx_a, nt * 4)); CK(cudaMalloc(&d_idx_b, nt * 4));
        CK(cudaMemcpyAsync(d_leaf, leaves_in, nt * 16, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(d_code_a, codes_in, nt * 8, cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
    } else {
        CK(cudaMalloc(&d_vert, nv * 12)); CK(cudaMalloc(&d_tri, nt * 12));
        CK(cudaMemcpyAsync(d_vert, vertices, nv * 12, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(d_tri, triangles, nt * 12, cudaMemcpyHostToDevice, st));
        CK(cudaMalloc(&d_bounds, 6 * 4)); CK(cudaMalloc(&d_hist, 64 * 8));
        // world coordinates: chroma/gpu/bvh.py:43-48
        uint32_t init[6] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u};
        CK(cudaMemcpyAsync(d_bounds, init, sizeof(init), cudaMemcpyHostToDevice, st));
        bounds_kernel<<<grid_for(nv, 256, c, 4), 256, 0, st>>>(d_vert, nv, d_bounds);
        c->launches++;
        CK(cudaMemcpyAsync(h_bounds, d_bounds, sizeof(h_bounds), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        float mn[3], mx[3], ext = 0.f;
        for (int k = 0; k < 3; k++) {
            mn[k] = ord2f(h_bounds[k]); mx[k] = ord2f(h_bounds[3 + k]);
            float e = mx[k] - mn[k];
            if (e > ext) ext = e;
            bvh->world_origin[k] = mn[k];
        }
        bvh->world_scale = ext / 65534.0f;
        if (!(bvh->world_scale > 0.f)) bvh->world_scale = 1.0f;  // degenerate mesh
        // leaves + Morton codes
        CK(cudaMalloc(&d_leaf, nt * 16)); CK(cudaMalloc(&d_leaf_sorted, nt * 16));
        CK(cudaMalloc(&d_code_a, nt * 8)); CK(cudaMalloc(&d_code_b, nt * 8));
        CK(cudaMalloc(&d_idx_a, nt * 4)); CK(cudaMalloc(&d_idx_b, nt * 4));
        make_leaves_kernel<<<grid, 256, 0, st>>>(nt, d_tri, d_vert, make_float3(mn[0], mn[1], mn[2]),
                                                  bvh->world_scale, d_leaf, d_code_a);
        c->launches++;
    }
    {
        iota_kernel<<<grid, 256, 0, st>>>(d_idx_a, nt);
        c->launches++;
        size_t tb = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, tb, d_code_a, d_code_b, d_idx_a, d_idx_b, (int)nt, 0, 48, st);
        {   // the same scratch serves the scans of the layer loop
            size_t sb = 0;
            cub::DeviceScan::ExclusiveSum(nullptr, sb, (uint32_t *)nullptr, (uint32_t *)nullptr, (int)nt, st);
            temp_cap = (tb > sb ? tb : sb) + (64u << 10);
        }
        CK(cudaMalloc(&d_temp, temp_cap));
        CK(cub::DeviceRadixSort::SortPairs(d_temp, tb, d_code_a, d_code_b, d_idx_a, d_idx_b, (int)nt, 0, 48, st));
        gather_nodes_kernel<<<grid, 256, 0, st>>>(d_leaf, d_idx_b, d_leaf_sorted, nt);
        c->launches += 2;
        CK(cudaFree(d_leaf)); d_leaf = nullptr;
        CK(cudaMalloc(&d_upper, upper_cap * 16));
        CK(cudaMalloc(&d_head, nt * 4)); CK(cudaMalloc(&d_gid, nt * 4)); CK(cudaMalloc(&d_gstart, (nt + 1) * 4));
        CK(cudaMalloc(&d_npar, nt * 4)); CK(cudaMalloc(&d_poff, nt * 4));
    }
    {
        // layer loop (grid.py:30-90)
        const uint4 *cur_nodes = d_leaf_sorted;
        unsigned long long *codes = d_code_b, *codes_next = d_code_a;
        uint64_t n = nt;
        while (n > 1) {
            unsigned g = grid_for(n, 256, c, 8);
            CK(cudaMemsetAsync(d_hist, 0, 64 * 8, st));
            diffbit_hist_kernel<<<g, 256, 0, st>>>(codes, n, d_hist);
            c->launches++;
            CK(cudaMemcpyAsync(h_hist, d_hist, sizeof(h_hist), cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            // nunique after shifting by s = 1 + #neighbour pairs whose highest differing bit >= s
            int shift = 0;
            for (;; shift++) {
                unsigned long long nunique = 1;
                for (int b = shift; b < 64; b++) nunique += h_hist[b];
                if (!((double)n / (double)nunique < (double)target_degree && nunique > 1)) break;
            }
            shift_heads_kernel<<<g, 256, 0, st>>>(codes, n, shift, d_head);
            c->launches++;
            size_t tb = temp_cap;
            CK(cub::DeviceScan::ExclusiveSum(d_temp, tb, d_head, d_gid, (int)n, st));
            uint32_t lh = 0, lg = 0;
            CK(cudaMemcpyAsync(&lh, d_head + n - 1, 4, cudaMemcpyDeviceToHost, st));
            CK(cudaMemcpyAsync(&lg, d_gid + n - 1, 4, cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            uint64_t ngroups = (uint64_t)lh + lg;
            if (shift) { apply_shift_kernel<<<g, 256, 0, st>>>(codes, n, shift); c->launches++; }
            group_start_kernel<<<g, 256, 0, st>>>(d_head, d_gid, n, d_gstart);
            unsigned gg = grid_for(ngroups, 256, c, 8);
            group_parents_kernel<<<gg, 256, 0, st>>>(d_gstart, ngroups, n, d_npar);
            c->launches += 3;
            tb = temp_cap;
            CK(cub::DeviceScan::ExclusiveSum(d_temp, tb, d_npar, d_poff, (int)ngroups, st));
            uint32_t ln = 0, lo = 0;
            CK(cudaMemcpyAsync(&ln, d_npar + ngroups - 1, 4, cudaMemcpyDeviceToHost, st));
            CK(cudaMemcpyAsync(&lo, d_poff + ngroups - 1, 4, cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            uint64_t nparents = (uint64_t)ln + lo;
            if (upper_used + nparents > upper_cap) {
                chroma_set_error("BVH upper-layer buffer overflow (%llu > %llu)",
                                 (unsigned long long)(upper_used + nparents), (unsigned long long)upper_cap);
                rc = CHROMA_ERR_OVERFLOW;
                goto done;
            }
            make_parents_kernel<<<gg, 256, 0, st>>>(d_gstart, d_poff, ngroups, n, cur_nodes, codes,
                                                    d_upper + upper_used, codes_next);
            c->launches++;
            layers.push_back({upper_used, nparents});
            cur_nodes = d_upper + upper_used;
            upper_used += nparents;
            unsigned long long *t = codes; codes = codes_next; codes_next = t;
            n = nparents;
        }
        // concatenate root-first (gpu/bvh.py:239-267) and collapse chains (gpu/bvh.py:114-130)
        uint64_t total = upper_used + nt;
        if (total >= (1ull << 28)) { chroma_set_error("BVH has too many nodes (%llu)", (unsigned long long)total); rc = CHROMA_ERR_OVERFLOW; goto done; }
        CK(cudaMalloc(&bvh->d_nodes, total * 16));
        bvh->nnodes = total;
        uint64_t pos = 0;
        std::vector<uint64_t> bounds;  // layer_bounds
        bounds.push_back(0);
        for (int li = (int)layers.size() - 1; li >= 0; li--) {
            uint64_t cnt = layers[li].count;
            uint64_t layer_end = pos + cnt;
            copy_and_offset_kernel<<<grid_for(cnt, 256, c, 8), 256, 0, st>>>(d_upper + layers[li].off, bvh->d_nodes + pos, cnt, (uint32_t)layer_end);
            c->launches++;
            pos = layer_end;
            bounds.push_back(pos);
        }
        CK(cudaMemcpyAsync(bvh->d_nodes + pos, d_leaf_sorted, nt * 16, cudaMemcpyDeviceToDevice, st));
        bounds.push_back(pos + nt);
        // all layers but the leaf layer, deepest first
        for (int li = (int)bounds.size() - 3; li >= 0; li--) {
            uint64_t s = bounds[li], e = bounds[li + 1];
            collapse_child_kernel<<<grid_for(e - s, 256, c, 8), 256, 0, st>>>(s, e, bvh->d_nodes);
            c->launches++;
        }
        CK(cudaStreamSynchronize(st));
        CK(cudaGetLastError());
        bvh->layer_offsets.assign(bounds.begin(), bounds.end() - 1);
    }
done:
    void *ptrs[] = {d_vert, d_tri, d_leaf, d_leaf_sorted, d_upper, d_code_a, d_code_b, d_hist, d_idx_a,
                    d_idx_b, d_head, d_gid, d_gstart, d_npar, d_poff, d_bounds, d_temp};
    for (void *p : ptrs)
        if (p) cudaFree(p);
    if (rc) {
        chroma_bvh_destroy(bvh);
        return rc;
    }
    *out = bvh;
    return 0;
}

extern "C" int chroma_bvh_build(chroma_ctx *c, const float *vertices, uint64_t nv,
                                const uint32_t *triangles, uint64_t nt, int target_degree,
                                chroma_bvh **out) {
    ARG_CHECK(vertices && triangles && nv > 0);
    return bvh_build_impl(c, vertices, nv, triangles, nt, nullptr, nullptr, nullptr, 0.f, target_degree, out);
}

extern "C" int chroma_bvh_build_from_leaves(chroma_ctx *c, const uint32_t *leaves, const uint64_t *codes,
                                            uint64_t nleaves, const float world_origin[3], float world_scale,
                                            int target_degree, chroma_bvh **out) {
    ARG_CHECK(leaves && codes && world_origin);
    return bvh_build_impl(c, nullptr, 0, nullptr, nleaves, leaves, codes, world_origin, world_scale,
                          target_degree, out);
}

extern "C" int chroma_bvh_destroy(chroma_bvh *b) {
    if (!b) return 0;
    cudaSetDevice(b->ctx->device);
    if (b->d_nodes) cudaFree(b->d_nodes);
    delete b;
    return 0;
}

extern "C" int chroma_bvh_info(chroma_bvh *b, uint64_t out[2], float world_origin[3], float *world_scale) {
    ARG_CHECK(b && out);
    out[0] = b->nnodes;
    out[1] = b->layer_offsets.size();
    if (world_origin) memcpy(world_origin, b->world_origin, 12);
    if (world_scale) *world_scale = b->world_scale;
    return 0;
}

extern "C" int chroma_bvh_get(chroma_bvh *b, uint32_t *nodes, uint64_t *layer_offsets) {
    ARG_CHECK(b);
    chroma_ctx *c = b->ctx;
    CTX_BIND(c);
    if (nodes) {
        CU_CHECK(cudaMemcpyAsync(nodes, b->d_nodes, b->nnodes * 16, cudaMemcpyDeviceToHost, c->stream));
        CU_CHECK(cudaStreamSynchronize(c->stream));
    }
    if (layer_offsets)
        for (size_t i = 0; i < b->layer_offsets.size(); i++) layer_offsets[i] = b->layer_offsets[i];
    return 0;
}

// ---- chroma/gpu/bvh.py helper API (host buffers in and out) --------------------------------
extern "C" int chroma_bvh_make_leaves(chroma_ctx *c, const float *vertices, uint64_t nv,
                                      const uint32_t *triangles, uint64_t nt, const float wo[3],
                                      float ws, uint32_t *leaf_nodes, uint64_t *morton_codes) {
    CTX_BIND(c);
    ARG_CHECK(vertices && triangles && wo && leaf_nodes && morton_codes && nv > 0 && nt > 0);
    int rc = 0;
    float *d_vert = nullptr; uint32_t *d_tri = nullptr; uint4 *d_leaf = nullptr; unsigned long long *d_code = nullptr;
    cudaStream_t st = c->stream;
    CK(cudaMalloc(&d_vert, nv * 12)); CK(cudaMalloc(&d_tri, nt * 12));
    CK(cudaMalloc(&d_leaf, nt * 16)); CK(cudaMalloc(&d_code, nt * 8));
    CK(cudaMemcpyAsync(d_vert, vertices, nv * 12, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_tri, triangles, nt * 12, cudaMemcpyHostToDevice, st));
    make_leaves_kernel<<<grid_for(nt, 256, c, 8), 256, 0, st>>>(nt, d_tri, d_vert, make_float3(wo[0], wo[1], wo[2]), ws, d_leaf, d_code);
    c->launches++;
    CK(cudaMemcpyAsync(leaf_nodes, d_leaf, nt * 16, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(morton_codes, d_code, nt * 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
done:
    if (d_vert) cudaFree(d_vert);
    if (d_tri) cudaFree(d_tri);
    if (d_leaf) cudaFree(d_leaf);
    if (d_code) cudaFree(d_code);
    return rc;
}

extern "C" int chroma_bvh_merge_nodes_detailed(chroma_ctx *c, const uint32_t *nodes, uint64_t nnodes,
                                               const int32_t *first_child, const int32_t *nchild,
                                               uint64_t nparent, uint32_t *parent_nodes) {
    CTX_BIND(c);
    ARG_CHECK(nodes && first_child && nchild && parent_nodes && nnodes > 0 && nparent > 0);
    int rc = 0;
    uint4 *d_nodes = nullptr, *d_par = nullptr; int *d_first = nullptr, *d_n = nullptr;
    cudaStream_t st = c->stream;
    CK(cudaMalloc(&d_nodes, nnodes * 16)); CK(cudaMalloc(&d_par, nparent * 16));
    CK(cudaMalloc(&d_first, nparent * 4)); CK(cudaMalloc(&d_n, nparent * 4));
    CK(cudaMemcpyAsync(d_nodes, nodes, nnodes * 16, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_first, first_child, nparent * 4, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_n, nchild, nparent * 4, cudaMemcpyHostToDevice, st));
    make_parents_detailed_kernel<<<grid_for(nparent, 256, c, 8), 256, 0, st>>>(nparent, d_nodes, d_par, d_first, d_n);
    c->launches++;
    CK(cudaMemcpyAsync(parent_nodes, d_par, nparent * 16, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
done:
    if (d_nodes) cudaFree(d_nodes);
    if (d_par) cudaFree(d_par);
    if (d_first) cudaFree(d_first);
    if (d_n) cudaFree(d_n);
    return rc;
}

extern "C" int chroma_bvh_collapse_chains(chroma_ctx *c, uint32_t *nodes, uint64_t nnodes,
                                          const uint64_t *layer_bounds, uint64_t nbounds) {
    CTX_BIND(c);
    ARG_CHECK(nodes && layer_bounds && nnodes > 0 && nbounds >= 2);
    int rc = 0;
    uint4 *d_nodes = nullptr;
    cudaStream_t st = c->stream;
    CK(cudaMalloc(&d_nodes, nnodes * 16));
    CK(cudaMemcpyAsync(d_nodes, nodes, nnodes * 16, cudaMemcpyHostToDevice, st));
    for (int64_t li = (int64_t)nbounds - 3; li >= 0; li--) {
        uint64_t s = layer_bounds[li], e = layer_bounds[li + 1];
        if (e <= s) continue;
        collapse_child_kernel<<<grid_for(e - s, 256, c, 8), 256, 0, st>>>(s, e, d_nodes);
        c->launches++;
    }
    CK(cudaMemcpyAsync(nodes, d_nodes, nnodes * 16, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaGetLastError());
done:
    if (d_nodes) cudaFree(d_nodes);
    return rc;
}
