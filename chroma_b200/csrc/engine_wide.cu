// engine_wide.cu — construction of the wide tree (wide.cuh) on the device and the kernels that
// search it: first-hit intersection and the fused persistent propagate kernel.
//
// Construction (all on the device, host only sequences the levels):
//   1. reference search rank of every leaf slot (tie break, see wide.cuh): subtree leaf counts
//      bottom up with arrival counters, per-child offsets, then one walk to the root per leaf
//   2. 63-bit Morton code of every triangle centroid (21 bits per axis on the world grid of the
//      reference BVH), cub radix sort of (code, slot)
//   3. triangle records, ranks and 16-bit padded boxes permuted into Morton order
//   4. groups top down, one level per launch: a range of the sorted codes is split at the
//      highest differing code bit (largest range first) until it has 8 parts or every part holds
//      <= leaf_max triangles; parts that small become leaves, the others are queued for the next level
//   5. boxes of inner records bottom up, level by level
// Replaces nothing in the reference (it has no such structure); the reference-format node array
// stays available as GPUGeometry.nodes and is still what the counting kernel walks.
#include <stdlib.h>
#include <mutex>
#include <cub/device/device_radix_sort.cuh>
#include "physics.cuh"
#include "wide.cuh"

static uint32_t env_u32w(const char *name, uint32_t dflt) {
    const char *s = getenv(name);
    if (!s || !*s) return dflt;
    return (uint32_t)strtoul(s, nullptr, 10);
}

// ---------------------------------------------------------------------------------------
// 1. reference search rank
// ---------------------------------------------------------------------------------------
#define NO_PARENT 0xFFFFFFFFu

__global__ void wk_parent_kernel(const uint4 *nodes, uint64_t nn, const uint32_t *reach, uint32_t *parent) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nn; i += stride) {
        if (!reach[i]) continue;
        uint32_t w = nodes[i].w, nchild = w >> 28, first = w & 0x0FFFFFFFu;
        for (uint32_t c = first; c < first + nchild && c < nn; c++) parent[c] = (uint32_t)i;
    }
}

// leaves start; the last child to arrive at a node sums the node's children and moves up
__global__ void wk_leafcount_kernel(const uint4 *nodes, uint64_t nn, const uint32_t *leaf_flag,
                                    const uint32_t *parent, uint32_t *arrive, uint32_t *cnt) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nn; i += stride) {
        if (!leaf_flag[i]) continue;
        cnt[i] = 1u;
        uint32_t p = parent[i];
        while (p != NO_PARENT) {
            __threadfence();
            uint32_t w = nodes[p].w, nchild = w >> 28, first = w & 0x0FFFFFFFu;
            uint32_t old = atomicAdd(arrive + p, 1u);
            if (old + 1u < nchild) break;
            __threadfence();
            uint32_t s = 0;
            for (uint32_t c = first; c < first + nchild; c++) s += __ldcg(cnt + c);
            cnt[p] = s;
            p = parent[p];
        }
    }
}

// offset of every child inside its parent's visiting order (mesh.h:64-111): the leaves of the
// group in index order, then the inner children from the last to the first
__global__ void wk_offsets_kernel(const uint4 *nodes, uint64_t nn, const uint32_t *reach,
                                  const uint32_t *cnt, uint32_t *off) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nn; i += stride) {
        if (!reach[i]) continue;
        uint32_t w = nodes[i].w, nchild = w >> 28, first = w & 0x0FFFFFFFu;
        if (nchild == 0) continue;
        uint32_t nleaf = 0;
        for (uint32_t c = first; c < first + nchild; c++)
            if ((nodes[c].w >> 28) == 0) off[c] = nleaf++;
        uint32_t acc = nleaf;
        for (uint32_t c = first + nchild; c-- > first;)
            if ((nodes[c].w >> 28) != 0) { off[c] = acc; acc += cnt[c]; }
    }
}

__global__ void wk_rank_kernel(uint64_t nn, const uint32_t *leaf_flag, const uint32_t *slot_of,
                               const uint32_t *parent, const uint32_t *off, uint32_t *rank) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nn; i += stride) {
        if (!leaf_flag[i]) continue;
        uint32_t r = 0, j = (uint32_t)i;
        while (j != 0u && j != NO_PARENT) {
            r += off[j];
            j = parent[j];
        }
        rank[slot_of[i]] = r;
    }
}

// ---------------------------------------------------------------------------------------
// 2./3. Morton codes, permutation
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t spread21(uint32_t v) {
    uint64_t x = v & 0x1FFFFFu;
    x = (x | x << 32) & 0x1F00000000FFFFull;
    x = (x | x << 16) & 0x1F0000FF0000FFull;
    x = (x | x << 8) & 0x100F00F00F00F00Full;
    x = (x | x << 4) & 0x10C30C30C30C30C3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void wk_codes_kernel(const float4 *tris, uint64_t n, float3 origin, float scale,
                                uint64_t *codes, uint32_t *vals) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const float inv = 32.0f / scale;   // 21 bits per axis = 16-bit world grid x 32
    for (uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; s < n; s += stride) {
        const float4 *p = tris + s * 3;
        float4 a = p[0], b = p[1], c = p[2];
        float cx = (a.x + a.w + b.z) * (1.0f / 3.0f), cy = (a.y + b.x + b.w) * (1.0f / 3.0f),
              cz = (a.z + b.y + c.x) * (1.0f / 3.0f);
        float qx = fminf(fmaxf((cx - origin.x) * inv, 0.0f), 2097151.0f);
        float qy = fminf(fmaxf((cy - origin.y) * inv, 0.0f), 2097151.0f);
        float qz = fminf(fmaxf((cz - origin.z) * inv, 0.0f), 2097151.0f);
        codes[s] = spread21((uint32_t)qx) | spread21((uint32_t)qy) << 1 | spread21((uint32_t)qz) << 2;
        vals[s] = (uint32_t)s;
    }
}

__device__ __forceinline__ uint32_t quant_lo(float v, float origin, float inv_scale) {
    float q = floorf((v - origin) * inv_scale) - 2.0f;   // two quanta of padding (bvh.cu:186-195 pads one;
                                                          // the second pays for the folded bias, wide.cuh)
    return (uint32_t)fminf(fmaxf(q, 0.0f), 65535.0f);
}
__device__ __forceinline__ uint32_t quant_hi(float v, float origin, float inv_scale) {
    float q = ceilf((v - origin) * inv_scale) + 2.0f;
    return (uint32_t)fminf(fmaxf(q, 0.0f), 65535.0f);
}

// triangle records, ranks and padded 16-bit boxes into Morton order; triangle id -> new slot
__global__ void wk_permute_kernel(const float4 *tris, const uint32_t *rank_old, const uint32_t *perm,
                                  uint64_t n, float3 origin, float scale, float4 *wtris,
                                  uint32_t *wrank, int32_t *wtri_slot, uint32_t *tbox) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const float is = 1.0f / scale;
    for (uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
        uint32_t s = perm[k];
        const float4 *p = tris + (size_t)s * 3;
        float4 a = p[0], b = p[1], c = p[2];
        float4 *q = wtris + k * 3;
        q[0] = a; q[1] = b; q[2] = c;
        wrank[k] = rank_old[s];
        wtri_slot[__float_as_uint(c.z)] = (int32_t)k;
        float lox = fminf(a.x, fminf(a.w, b.z)), hix = fmaxf(a.x, fmaxf(a.w, b.z));
        float loy = fminf(a.y, fminf(b.x, b.w)), hiy = fmaxf(a.y, fmaxf(b.x, b.w));
        float loz = fminf(a.z, fminf(b.y, c.x)), hiz = fmaxf(a.z, fmaxf(b.y, c.x));
        tbox[3 * k + 0] = quant_lo(lox, origin.x, is) | quant_hi(hix, origin.x, is) << 16;
        tbox[3 * k + 1] = quant_lo(loy, origin.y, is) | quant_hi(hiy, origin.y, is) << 16;
        tbox[3 * k + 2] = quant_lo(loz, origin.z, is) | quant_hi(hiz, origin.z, is) << 16;
    }
}

// ---------------------------------------------------------------------------------------
// 4. groups, one level per launch
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t box_union(uint32_t a, uint32_t b) {
    uint32_t lo = min(a & 0xFFFFu, b & 0xFFFFu), hi = max(a >> 16, b >> 16);
    return lo | hi << 16;
}
// after the permutation the reference-format engine nodes point at the Morton-order slots too,
// so one triangle array serves both trees
__global__ void wk_remap_leaves_kernel(const uint4 *nodes_ref, uint64_t nn, const uint32_t *leaf_flag,
                                       const int32_t *wtri_slot, uint4 *nodes) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nn; i += stride)
        if (leaf_flag[i]) nodes[i].w = (uint32_t)wtri_slot[nodes_ref[i].w];
}

#ifndef WIDE_SAH_DEFAULT
#define WIDE_SAH_DEFAULT 0u   // parts up to this many triangles are split by area cost (sah_split); 0 = Morton prefix only
#endif
#define BOX_NONE 0x0000FFFFu  // lo = 65535, hi = 0: neutral element of box_union

// first index of the upper part when [b,e) is cut at the highest bit in which its codes differ
__device__ __forceinline__ uint32_t wide_split(const uint64_t *codes, uint32_t b, uint32_t e) {
    uint64_t first = codes[b], last = codes[e - 1];
    if (first == last) return b + ((e - b) >> 1);
    int prefix = __clzll((long long)(first ^ last));
    uint32_t lo = b, hi = e - 1;   // codes[lo] agrees with `first` beyond `prefix` bits, codes[hi] does not
    while (hi - lo > 1) {
        uint32_t mid = lo + ((hi - lo) >> 1);
        if (__clzll((long long)(first ^ codes[mid])) > prefix) lo = mid;
        else hi = mid;
    }
    return hi;
}

// Split of a part no larger than `sah_max` by surface-area cost instead of by Morton prefix
// (CHROMA_WIDE_SAH, r2h): the part stays a contiguous range of the Morton order; up to 32 evenly
// spaced split positions are tried (every position when the part has <= 32 triangles) and the one
// with the least  area(left) * n(left) + area(right) * n(right)  is taken.  The search result does
// not depend on the shape of the tree (ties go by the reference visiting rank), only its cost does.
__device__ __forceinline__ float box_area(uint32_t x, uint32_t y, uint32_t z) {
    float dx = (float)((x >> 16) - (x & 0xFFFFu)), dy = (float)((y >> 16) - (y & 0xFFFFu)),
          dz = (float)((z >> 16) - (z & 0xFFFFu));
    return dx * dy + dy * dz + dz * dx;
}

__device__ uint32_t sah_split(const uint32_t *__restrict__ tbox, uint32_t b, uint32_t e) {
    const uint32_t n = e - b;
    if (n <= 2) return b + (n >> 1);
    const uint32_t K = n < 32u ? n : 32u;
    uint32_t bx[32], by[32], bz[32];
    for (uint32_t k = 0; k < K; k++) {
        uint32_t s = b + (uint32_t)((uint64_t)k * n / K), t = b + (uint32_t)((uint64_t)(k + 1) * n / K);
        uint32_t x = BOX_NONE, y = BOX_NONE, z = BOX_NONE;
        for (uint32_t i = s; i < t; i++) {
            x = box_union(x, tbox[3 * (size_t)i]);
            y = box_union(y, tbox[3 * (size_t)i + 1]);
            z = box_union(z, tbox[3 * (size_t)i + 2]);
        }
        bx[k] = x; by[k] = y; bz[k] = z;
    }
    float right[32];
    {
        uint32_t x = BOX_NONE, y = BOX_NONE, z = BOX_NONE;
        for (uint32_t k = K - 1; k >= 1; k--) {
            x = box_union(x, bx[k]); y = box_union(y, by[k]); z = box_union(z, bz[k]);
            right[k] = box_area(x, y, z);
        }
    }
    uint32_t x = BOX_NONE, y = BOX_NONE, z = BOX_NONE, best = K >> 1;
    float best_cost = 3.0e38f;
    for (uint32_t k = 1; k < K; k++) {   // split before bucket k
        x = box_union(x, bx[k - 1]); y = box_union(y, by[k - 1]); z = box_union(z, bz[k - 1]);
        const uint32_t nl = (uint32_t)((uint64_t)k * n / K);
        const float cost = box_area(x, y, z) * (float)nl + right[k] * (float)(n - nl);
        if (cost < best_cost) { best_cost = cost; best = k; }
    }
    return b + (uint32_t)((uint64_t)best * n / K);
}

// item = (begin, end, parent record index or 0xFFFFFFFF for the root, unused)
__global__ void wk_level_kernel(const uint64_t *codes, const uint32_t *tbox, const uint4 *items,
                                uint32_t nitems, uint32_t leaf_max, uint32_t fat_max, uint32_t sah_max, uint4 *groups,
                                uint32_t group_cap,
                                uint32_t *counters /* [0] groups, [1] items out, [2] error, [3] leaf records */,
                                uint4 *items_out, uint32_t items_cap) {
    uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= nitems) return;
    uint4 it = items[id];
    uint32_t cb[8], ce[8];
    int nc = 1;
    cb[0] = it.x; ce[0] = it.y;
    while (nc < 8) {
        int best = -1;
        uint32_t bs = leaf_max;
        for (int j = 0; j < nc; j++) {
            uint32_t sz = ce[j] - cb[j];
            if (sz > bs) { bs = sz; best = j; }
        }
        if (best < 0) break;
        uint32_t m = bs <= sah_max ? sah_split(tbox, cb[best], ce[best]) : wide_split(codes, cb[best], ce[best]);
        for (int j = nc; j > best + 1; j--) { cb[j] = cb[j - 1]; ce[j] = ce[j - 1]; }
        cb[best + 1] = m; ce[best + 1] = ce[best]; ce[best] = m;
        nc++;
    }
    uint32_t g = atomicAdd(counters + 0, 1u);
    if (g >= group_cap) { atomicOr(counters + 2, 1u); return; }
    if (it.z != 0xFFFFFFFFu) groups[it.z].w = g;   // the parent record now points here
    uint4 *rec = groups + (size_t)g * 8;
    uint32_t nleaf = 0;
    for (int j = 0; j < 8; j++) {
        uint4 r = make_uint4(BOX_NONE, BOX_NONE, BOX_NONE, WIDE_EMPTY);
        if (j < nc) {
            uint32_t sz = ce[j] - cb[j];
            // a part just over leaf_max that found no free record to be split into stays one
            // (longer) leaf rather than becoming a two-leaf group one level down
            if (sz <= fat_max) {
                nleaf++;
                for (uint32_t k = cb[j]; k < ce[j]; k++) {
                    r.x = box_union(r.x, tbox[3 * (size_t)k]);
                    r.y = box_union(r.y, tbox[3 * (size_t)k + 1]);
                    r.z = box_union(r.z, tbox[3 * (size_t)k + 2]);
                }
                r.w = WIDE_LEAF | (sz - 1u) << 28 | cb[j];
            } else {
                uint32_t o = atomicAdd(counters + 1, 1u);
                if (o >= items_cap) { atomicOr(counters + 2, 2u); }
                else items_out[o] = make_uint4(cb[j], ce[j], g * 8u + (uint32_t)j, 0u);
                r.w = 0u;   // child group index, written by the child at the next level
            }
        }
        rec[j] = r;
    }
    atomicAdd(counters + 3, nleaf);
}

// 5. boxes of the inner records of groups [gbegin, gend): union of the child group's records
__global__ void wk_refit_kernel(uint4 *groups, uint32_t gbegin, uint32_t gend) {
    uint64_t id = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t rec = (uint64_t)gbegin * 8 + id;
    if (rec >= (uint64_t)gend * 8) return;
    uint4 r = groups[rec];
    if (r.w == WIDE_EMPTY || (r.w & WIDE_LEAF)) return;
    const uint4 *c = groups + (size_t)r.w * 8;
    uint32_t x = BOX_NONE, y = BOX_NONE, z = BOX_NONE;
    for (int j = 0; j < 8; j++) {
        uint4 cr = c[j];
        if (cr.w == WIDE_EMPTY) break;
        x = box_union(x, cr.x); y = box_union(y, cr.y); z = box_union(z, cr.z);
    }
    r.x = x; r.y = y; r.z = z;
    groups[rec] = r;
}

struct WideTemps {
    uint32_t *parent = nullptr, *arrive = nullptr, *cnt = nullptr, *off = nullptr, *rank_old = nullptr;
    uint64_t *codes = nullptr, *codes2 = nullptr;
    uint32_t *vals = nullptr, *perm = nullptr, *tbox = nullptr, *counters = nullptr;
    uint4 *items[2] = {nullptr, nullptr}, *groups_big = nullptr;
    void *temp = nullptr;
    template <typename T> static void drop(T *&p) { if (p) cudaFree(p); p = nullptr; }
    void release() {
        drop(parent); drop(arrive); drop(cnt); drop(off); drop(rank_old); drop(codes); drop(codes2);
        drop(vals); drop(perm); drop(tbox); drop(counters); drop(items[0]); drop(items[1]);
        drop(groups_big); drop(temp);
    }
};

static int wide_build_impl(chroma_geometry *g, const uint32_t *d_reach, const uint32_t *d_leaf_flag,
                           const uint32_t *d_slot_of, uint32_t leaf_max, WideTemps &T) {
    chroma_ctx *c = g->ctx;
    const uint64_t nn = g->nnodes, n = g->nslots;
    WideView &W = g->wide;
    const unsigned grid_nn = grid_for(nn, 256, c, 8), grid_n = grid_for(n, 256, c, 8);
    // ---- 1. reference search rank ----
    CU_CHECK(cudaMalloc(&T.parent, nn * 4)); CU_CHECK(cudaMalloc(&T.arrive, nn * 4));
    CU_CHECK(cudaMalloc(&T.cnt, nn * 4)); CU_CHECK(cudaMalloc(&T.off, nn * 4));
    CU_CHECK(cudaMalloc(&T.rank_old, n * 4));
    CU_CHECK(cudaMemsetAsync(T.parent, 0xFF, nn * 4, c->stream));
    CU_CHECK(cudaMemsetAsync(T.arrive, 0, nn * 4, c->stream));
    CU_CHECK(cudaMemsetAsync(T.cnt, 0, nn * 4, c->stream));
    CU_CHECK(cudaMemsetAsync(T.off, 0, nn * 4, c->stream));
    wk_parent_kernel<<<grid_nn, 256, 0, c->stream>>>(g->d_nodes_ref, nn, d_reach, T.parent);
    wk_leafcount_kernel<<<grid_nn, 256, 0, c->stream>>>(g->d_nodes_ref, nn, d_leaf_flag, T.parent, T.arrive, T.cnt);
    wk_offsets_kernel<<<grid_nn, 256, 0, c->stream>>>(g->d_nodes_ref, nn, d_reach, T.cnt, T.off);
    wk_rank_kernel<<<grid_nn, 256, 0, c->stream>>>(nn, d_leaf_flag, d_slot_of, T.parent, T.off, T.rank_old);
    c->launches += 4;
    CU_CHECK(cudaGetLastError());
    CU_CHECK(cudaStreamSynchronize(c->stream));
    WideTemps::drop(T.parent); WideTemps::drop(T.arrive); WideTemps::drop(T.cnt); WideTemps::drop(T.off);

    // ---- 2. Morton codes + sort ----
    CU_CHECK(cudaMalloc(&T.codes, n * 8)); CU_CHECK(cudaMalloc(&T.codes2, n * 8));
    CU_CHECK(cudaMalloc(&T.vals, n * 4)); CU_CHECK(cudaMalloc(&T.perm, n * 4));
    wk_codes_kernel<<<grid_n, 256, 0, c->stream>>>(g->d_tris, n, W.world_origin, W.world_scale, T.codes, T.vals);
    c->launches++;
    size_t tb = 0;
    CU_CHECK(cub::DeviceRadixSort::SortPairs(nullptr, tb, T.codes, T.codes2, T.vals, T.perm, (int)n, 0, 63, c->stream));
    CU_CHECK(cudaMalloc(&T.temp, tb ? tb : 16));
    CU_CHECK(cub::DeviceRadixSort::SortPairs(T.temp, tb, T.codes, T.codes2, T.vals, T.perm, (int)n, 0, 63, c->stream));
    c->launches++;
    CU_CHECK(cudaStreamSynchronize(c->stream));
    WideTemps::drop(T.temp); WideTemps::drop(T.codes); WideTemps::drop(T.vals);

    // ---- 3. permute ----
    CU_CHECK(cudaMalloc(&g->d_wtris, n * 3 * sizeof(float4)));
    CU_CHECK(cudaMalloc(&g->d_wrank, n * 4));
    CU_CHECK(cudaMalloc(&g->d_wtri_slot, g->ntriangles * 4));
    CU_CHECK(cudaMalloc(&T.tbox, n * 12));
    CU_CHECK(cudaMemsetAsync(g->d_wtri_slot, 0xFF, g->ntriangles * 4, c->stream));
    wk_permute_kernel<<<grid_n, 256, 0, c->stream>>>(g->d_tris, T.rank_old, T.perm, n, W.world_origin,
                                                      W.world_scale, g->d_wtris, g->d_wrank,
                                                      g->d_wtri_slot, T.tbox);
    c->launches++;
    CU_CHECK(cudaStreamSynchronize(c->stream));
    WideTemps::drop(T.rank_old); WideTemps::drop(T.perm);

    // ---- 4. groups ----
    // every queued part holds > leaf_max triangles and the parts of one level are disjoint
    const uint32_t items_cap = (uint32_t)(n / (leaf_max + 1) + 16);
    const uint32_t group_cap = (uint32_t)(n / (leaf_max + 1) * 2 + 16);
    CU_CHECK(cudaMalloc(&T.items[0], (size_t)items_cap * sizeof(uint4)));
    CU_CHECK(cudaMalloc(&T.items[1], (size_t)items_cap * sizeof(uint4)));
    CU_CHECK(cudaMalloc(&T.groups_big, (size_t)group_cap * 8 * sizeof(uint4)));
    CU_CHECK(cudaMalloc(&T.counters, 16));
    CU_CHECK(cudaMemsetAsync(T.counters, 0, 16, c->stream));
    uint4 root_item = make_uint4(0u, (uint32_t)n, 0xFFFFFFFFu, 0u);
    CU_CHECK(cudaMemcpyAsync(T.items[0], &root_item, sizeof(uint4), cudaMemcpyHostToDevice, c->stream));
    CU_CHECK(cudaStreamSynchronize(c->stream));
    uint32_t fat_max = env_u32w("CHROMA_WIDE_FAT", 0);
    if (fat_max < leaf_max) fat_max = leaf_max;
    if (fat_max > 8) fat_max = 8;
    const uint32_t sah_max = env_u32w("CHROMA_WIDE_SAH", WIDE_SAH_DEFAULT);
    std::vector<uint32_t> level_end;   // number of groups after each level
    uint32_t nitems = 1, ngroups = 0;
    int cur = 0;
    while (nitems > 0) {
        if (level_end.size() >= 128) {
            chroma_set_error("wide tree build: more than 128 levels");
            return CHROMA_ERR_OVERFLOW;
        }
        CU_CHECK(cudaMemsetAsync(T.counters + 1, 0, 4, c->stream));
        wk_level_kernel<<<(nitems + 127) / 128, 128, 0, c->stream>>>(
            T.codes2, T.tbox, T.items[cur], nitems, leaf_max, fat_max, sah_max, T.groups_big, group_cap, T.counters,
            T.items[cur ^ 1], items_cap);
        c->launches++;
        uint32_t h[4];
        CU_CHECK(cudaMemcpyAsync(h, T.counters, 16, cudaMemcpyDeviceToHost, c->stream));
        CU_CHECK(cudaStreamSynchronize(c->stream));
        if (h[2]) {
            chroma_set_error("wide tree build: capacity exceeded (%u)", h[2]);
            return CHROMA_ERR_OVERFLOW;
        }
        ngroups = h[0];
        nitems = h[1];
        g->wide_leaves = h[3];
        level_end.push_back(ngroups);
        cur ^= 1;
    }
    // ---- 5. refit, deepest level first ----
    for (int l = (int)level_end.size() - 1; l >= 0; l--) {
        uint32_t gb = l ? level_end[l - 1] : 0u, ge = level_end[l];
        uint64_t nrec = (uint64_t)(ge - gb) * 8;
        if (nrec == 0) continue;
        wk_refit_kernel<<<(unsigned)((nrec + 255) / 256), 256, 0, c->stream>>>(T.groups_big, gb, ge);
        c->launches++;
    }
    CU_CHECK(cudaGetLastError());
    CU_CHECK(cudaMalloc(&g->d_wgroups, (size_t)ngroups * 8 * sizeof(uint4)));
    CU_CHECK(cudaMemcpyAsync(g->d_wgroups, T.groups_big, (size_t)ngroups * 8 * sizeof(uint4),
                             cudaMemcpyDeviceToDevice, c->stream));
    CU_CHECK(cudaStreamSynchronize(c->stream));
    g->bytes += (size_t)ngroups * 128 + n * 52 + g->ntriangles * 4;
    W.groups = g->d_wgroups;
    W.tris = g->d_wtris;
    W.rank = g->d_wrank;
    W.ngroups = ngroups;
    W.nslots = (uint32_t)n;
    g->wide_levels = (uint32_t)level_end.size();
    // one triangle array for both trees: the reference-format nodes now carry Morton-order slots
    wk_remap_leaves_kernel<<<grid_nn, 256, 0, c->stream>>>(g->d_nodes_ref, nn, d_leaf_flag, g->d_wtri_slot, g->d_nodes);
    c->launches++;
    CU_CHECK(cudaGetLastError());
    CU_CHECK(cudaStreamSynchronize(c->stream));
    cudaFree(g->d_tris); cudaFree(g->d_tri_slot);
    g->bytes -= n * 3 * sizeof(float4) + g->ntriangles * 4;
    g->d_tris = g->d_wtris; g->d_tri_slot = g->d_wtri_slot;
    g->d_wtris = nullptr; g->d_wtri_slot = nullptr;
    g->view.tris = g->d_tris;
    g->view.tri_slot = g->d_tri_slot;
    g->wview = g->view;
    return 0;
}

// ---------------------------------------------------------------------------------------
// Generic form of steps 2, 4, 5 for the two-level tree (engine_inst.cu): an 8-wide tree over `n`
// items given by their Morton codes and padded 16-bit boxes ([n][3] words lo | hi << 16).  Leaf
// records refer to positions in the SORTED order; out->perm[k] is the input index of sorted item k.
// ---------------------------------------------------------------------------------------
__global__ void wk_iota_kernel(uint32_t *a, uint64_t n) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) a[i] = (uint32_t)i;
}
__global__ void wk_gather_boxes_kernel(const uint32_t *boxes, const uint32_t *perm, uint64_t n, uint32_t *tbox) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
        const uint32_t s = perm[k];
        tbox[3 * k] = boxes[3 * (size_t)s]; tbox[3 * k + 1] = boxes[3 * (size_t)s + 1]; tbox[3 * k + 2] = boxes[3 * (size_t)s + 2];
    }
}

static int wide_tree_impl(chroma_ctx *c, uint64_t n, const uint64_t *d_codes, const uint32_t *d_boxes,
                          uint32_t leaf_max, WideTreeOut *out, WideTemps &T) {
    const unsigned grid_n = grid_for(n, 256, c, 8);
    CU_CHECK(cudaMalloc(&T.codes, n * 8)); CU_CHECK(cudaMalloc(&T.codes2, n * 8));
    CU_CHECK(cudaMalloc(&T.vals, n * 4)); CU_CHECK(cudaMalloc(&T.perm, n * 4));
    CU_CHECK(cudaMemcpyAsync(T.codes, d_codes, n * 8, cudaMemcpyDeviceToDevice, c->stream));
    wk_iota_kernel<<<grid_n, 256, 0, c->stream>>>(T.vals, n);
    c->launches++;
    size_t tb = 0;
    CU_CHECK(cub::DeviceRadixSort::SortPairs(nullptr, tb, T.codes, T.codes2, T.vals, T.perm, (int)n, 0, 63, c->stream));
    CU_CHECK(cudaMalloc(&T.temp, tb ? tb : 16));
    CU_CHECK(cub::DeviceRadixSort::SortPairs(T.temp, tb, T.codes, T.codes2, T.vals, T.perm, (int)n, 0, 63, c->stream));
    c->launches++;
    CU_CHECK(cudaMalloc(&T.tbox, n * 12));
    wk_gather_boxes_kernel<<<grid_n, 256, 0, c->stream>>>(d_boxes, T.perm, n, T.tbox);
    c->launches++;
    const uint32_t items_cap = (uint32_t)(n / (leaf_max + 1) + 16);
    const uint32_t group_cap = (uint32_t)(n / (leaf_max + 1) * 2 + 16);
    CU_CHECK(cudaMalloc(&T.items[0], (size_t)items_cap * sizeof(uint4)));
    CU_CHECK(cudaMalloc(&T.items[1], (size_t)items_cap * sizeof(uint4)));
    CU_CHECK(cudaMalloc(&T.groups_big, (size_t)group_cap * 8 * sizeof(uint4)));
    CU_CHECK(cudaMalloc(&T.counters, 16));
    CU_CHECK(cudaMemsetAsync(T.counters, 0, 16, c->stream));
    uint4 root_item = make_uint4(0u, (uint32_t)n, 0xFFFFFFFFu, 0u);
    CU_CHECK(cudaMemcpyAsync(T.items[0], &root_item, sizeof(uint4), cudaMemcpyHostToDevice, c->stream));
    CU_CHECK(cudaStreamSynchronize(c->stream));
    const uint32_t sah_max = env_u32w("CHROMA_WIDE_SAH", WIDE_SAH_DEFAULT);
    std::vector<uint32_t> level_end;
    uint32_t nitems = 1, ngroups = 0, nleaves = 0;
    int cur = 0;
    while (nitems > 0) {
        if (level_end.size() >= 128) {
            chroma_set_error("wide tree build: more than 128 levels");
            return CHROMA_ERR_OVERFLOW;
        }
        CU_CHECK(cudaMemsetAsync(T.counters + 1, 0, 4, c->stream));
        wk_level_kernel<<<(nitems + 127) / 128, 128, 0, c->stream>>>(
            T.codes2, T.tbox, T.items[cur], nitems, leaf_max, leaf_max, sah_max, T.groups_big, group_cap, T.counters,
            T.items[cur ^ 1], items_cap);
        c->launches++;
        uint32_t h[4];
        CU_CHECK(cudaMemcpyAsync(h, T.counters, 16, cudaMemcpyDeviceToHost, c->stream));
        CU_CHECK(cudaStreamSynchronize(c->stream));
        if (h[2]) {
            chroma_set_error("wide tree build: capacity exceeded (%u)", h[2]);
            return CHROMA_ERR_OVERFLOW;
        }
        ngroups = h[0]; nitems = h[1]; nleaves = h[3];
        level_end.push_back(ngroups);
        cur ^= 1;
    }
    for (int l = (int)level_end.size() - 1; l >= 0; l--) {
        uint32_t gb = l ? level_end[l - 1] : 0u, ge = level_end[l];
        uint64_t nrec = (uint64_t)(ge - gb) * 8;
        if (nrec == 0) continue;
        wk_refit_kernel<<<(unsigned)((nrec + 255) / 256), 256, 0, c->stream>>>(T.groups_big, gb, ge);
        c->launches++;
    }
    CU_CHECK(cudaGetLastError());
    CU_CHECK(cudaMalloc(&out->groups, (size_t)ngroups * 8 * sizeof(uint4)));
    CU_CHECK(cudaMemcpyAsync(out->groups, T.groups_big, (size_t)ngroups * 8 * sizeof(uint4),
                             cudaMemcpyDeviceToDevice, c->stream));
    CU_CHECK(cudaStreamSynchronize(c->stream));
    out->ngroups = ngroups;
    out->levels = (uint32_t)level_end.size();
    out->leaves = nleaves;
    out->perm = T.perm;
    T.perm = nullptr;   // handed to the caller
    return 0;
}

int wide_tree_from_boxes(chroma_ctx *c, uint64_t n, const uint64_t *d_codes, const uint32_t *d_boxes,
                         uint32_t leaf_max, WideTreeOut *out) {
    *out = WideTreeOut{};
    if (n == 0) return 0;
    WideTemps T;
    int rc = wide_tree_impl(c, n, d_codes, d_boxes, leaf_max, out, T);
    T.release();
    if (rc) {
        if (out->groups) cudaFree(out->groups);
        if (out->perm) cudaFree(out->perm);
        *out = WideTreeOut{};
    }
    return rc;
}

int chroma_wide_build(chroma_geometry *g, const uint32_t *d_reach, const uint32_t *d_leaf_flag,
                      const uint32_t *d_slot_of) {
    chroma_ctx *c = g->ctx;
    WideView &W = g->wide;
    W = WideView{};
    W.world_origin = g->view.world_origin;
    W.world_scale = g->view.world_scale;
    g->wview = g->view;
    g->wide_levels = 0;
    if (env_u32w("CHROMA_WIDE", 1) == 0 || g->nslots == 0) return 0;
    // a leaf at the root is never tested by the reference (mesh.h:60-64 pushes its zero children)
    uint4 root;
    CU_CHECK(cudaMemcpyAsync(&root, g->d_nodes_ref, sizeof(uint4), cudaMemcpyDeviceToHost, c->stream));
    CU_CHECK(cudaStreamSynchronize(c->stream));
    if ((root.w >> 28) == 0) return 0;
    uint32_t leaf_max = env_u32w("CHROMA_WIDE_LEAF", 3);
    if (leaf_max < 1) leaf_max = 1;
    if (leaf_max > 8) leaf_max = 8;
    WideTemps T;
    int rc = wide_build_impl(g, d_reach, d_leaf_flag, d_slot_of, leaf_max, T);
    T.release();
    // the product library has no other search kernel: failing to build the tree fails the call.  A/B
    // builds can fall back to the reference-format tree unless CHROMA_STRICT=1
    if (rc && (c->knobs.strict || !CHROMA_AB_KERNELS)) return rc;
    if (rc) {
        // the engine stays usable on the reference-format tree (about half the rate); the caller sees
        // it in chroma_geometry_tree (out[0] == 0) and on stderr
        fprintf(stderr, "chroma_b200: wide tree not built (%s); searching the reference tree\n", chroma_last_error());
        cudaGetLastError();
        void *own[] = {g->d_wgroups, g->d_wtris, g->d_wrank, g->d_wtri_slot};
        for (void *p : own) if (p) cudaFree(p);
        g->d_wgroups = nullptr; g->d_wtris = nullptr; g->d_wrank = nullptr; g->d_wtri_slot = nullptr;
        W = WideView{};
        W.world_origin = g->view.world_origin;
        W.world_scale = g->view.world_scale;
        g->wview = g->view;
        g->wide_levels = 0;
    }
    return 0;
}

// ---------------------------------------------------------------------------------------
// first-hit intersection on the wide tree (mesh.h:123-151 semantics, plus the triangle index)
// ---------------------------------------------------------------------------------------
#define WISECT_THREADS 128

struct WIsectParams {
    unsigned long long *head;
    int refill, iters, leaf_thresh;
};

template <bool COUNT>
__global__ void __launch_bounds__(WISECT_THREADS)
wide_intersect_kernel(WideView w, const int32_t *__restrict__ tri_slot, uint32_t ntriangles, uint64_t n,
                      const float *__restrict__ pos, const float *__restrict__ dir,
                      const int32_t *__restrict__ last_hit, int32_t *__restrict__ out_tri,
                      float *__restrict__ out_dist, uint32_t *__restrict__ out_counts, uint32_t *status,
                      WIsectParams prm) {
    extern __shared__ float4 smem_raw[];
    uint2 *s_stack = (uint2 *)smem_raw;
    uint2 l_stack[WIDE_STACK - WIDE_SSTACK];
    const WStackRef stk = wstk_make(s_stack + threadIdx.x, l_stack, WISECT_THREADS);
    WSpread spr;
    {
        unsigned char *sb = (unsigned char *)(s_stack + WIDE_SSTACK * WISECT_THREADS) + (threadIdx.x >> 5) * WSPREAD_BYTES_PER_WARP;
        spr.res = (float *)sb; spr.own = sb + WSPREAD_PAIRS * 4;
    }
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    bool have = false, exhausted = false;
    unsigned overflow = 0u;
    float3 o = f3(0, 0, 0), d = f3(0, 0, 1);
    uint64_t id = 0;
    uint32_t ng = 0, nt = 0;
    WTrav tv;
    tv.cur = WIDE_NONE; tv.sp = -1; tv.hit = -1; tv.dist = 0.f; tv.last_slot = -1;
    while (true) {
        const bool work = have && !wtrav_done(tv);
        const unsigned mW = __ballot_sync(0xffffffffu, work);
        const unsigned mH = __ballot_sync(0xffffffffu, have);
        if (mH == 0u && exhausted) break;
        const int nidle = __popc(mH & ~mW) + (exhausted ? 0 : 32 - __popc(mH));
        if (mW == 0u || nidle >= prm.refill) {
            if (have && !work) {
                int tri = -1;
                if (tv.hit >= 0) tri = (int)__float_as_uint(__ldg(w.tris + (size_t)tv.hit * 3 + 2).z);
                out_tri[id] = tri;
                out_dist[id] = tv.hit >= 0 ? tv.dist : -1.0f;
                if (COUNT) { out_counts[2 * id] = ng; out_counts[2 * id + 1] = nt; }
                have = false;
            }
            if (!exhausted) {
                const unsigned hv = __ballot_sync(0xffffffffu, have);
                const int nfree = 32 - __popc(hv);
                if (nfree > 0) {
                    unsigned long long base = 0;
                    if (lane == 0) base = atomicAdd(prm.head, (unsigned long long)nfree);
                    base = __shfl_sync(0xffffffffu, base, 0);
                    if (base + (unsigned long long)nfree >= n) exhausted = true;
                    if (!have) {
                        unsigned long long i = base + (unsigned long long)__popc(~hv & lt_mask);
                        if (i < n) {
                            id = i;
                            o = f3(pos[3 * id], pos[3 * id + 1], pos[3 * id + 2]);
                            d = f3(dir[3 * id], dir[3 * id + 1], dir[3 * id + 2]);
                            // mesh.h:141  direction /= norm(direction)  (sqrt.approx, div.approx)
                            float nrm = sqrtf(ref_dot(d, d));
                            d = f3(__fdividef(d.x, nrm), __fdividef(d.y, nrm), __fdividef(d.z, nrm));
                            int last_slot = -1;
                            if (last_hit) {
                                int lt = last_hit[id];
                                if (lt >= 0 && (uint32_t)lt < ntriangles) last_slot = tri_slot[lt];
                            }
                            wtrav_begin(tv, w, o, d, last_slot);
                            if (wide_origin_unsearchable(o, WIDE_GUARD_QUANTA * w.world_scale)) tv.cur = WIDE_NONE;   // no hit
                            ng = nt = 0;
                            have = true;
                        }
                    }
                }
            }
            continue;
        }
        for (int it = 0; it < prm.iters; it++)
            wtrav_round<COUNT>(tv, have, w, o, d, stk, prm.leaf_thresh, overflow, ng, nt, spr);
    }
    if (overflow) atomicOr(status, 1u);
}

int chroma_wide_intersect(chroma_geometry *g, uint64_t n, const float *pos, const float *dir,
                          const int32_t *last_hit, int32_t *out_tri, float *out_dist,
                          uint32_t *out_counts) {
    chroma_ctx *c = g->ctx;
    if (n == 0) return 0;
    if (g->instanced) return chroma_inst_intersect(g, n, pos, dir, last_hit, out_tri, out_dist, out_counts);
    uint32_t *status = c->d_scratch + 2;
    CU_CHECK(cudaMemsetAsync(status, 0, 4, c->stream));
    WIsectParams prm;
    prm.head = (unsigned long long *)(c->d_scratch + 10);
    CU_CHECK(cudaMemsetAsync(prm.head, 0, 8, c->stream));
    prm.refill = c->knobs.isect_refill;   // no physics to batch: refill at once (r02b)
    prm.iters = c->knobs.isect_iters;
    prm.leaf_thresh = c->knobs.leaf_thresh;
    size_t smem = (size_t)WIDE_SSTACK * WISECT_THREADS * sizeof(uint2) + (WIDE_LEAF_SPREAD ? (WISECT_THREADS / 32) * WSPREAD_BYTES_PER_WARP : 0);
    int occ = 0;
    if (out_counts) {
        CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, wide_intersect_kernel<true>, WISECT_THREADS, smem));
        if (occ < 1) occ = 1;
        unsigned grid = grid_for(n, WISECT_THREADS, c, (unsigned)occ);
        wide_intersect_kernel<true><<<grid, WISECT_THREADS, smem, c->stream>>>(
            g->wide, g->d_tri_slot, (uint32_t)g->ntriangles, n, pos, dir, last_hit, out_tri, out_dist,
            out_counts, status, prm);
    } else {
        CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, wide_intersect_kernel<false>, WISECT_THREADS, smem));
        if (occ < 1) occ = 1;
        unsigned grid = grid_for(n, WISECT_THREADS, c, (unsigned)occ);
        wide_intersect_kernel<false><<<grid, WISECT_THREADS, smem, c->stream>>>(
            g->wide, g->d_tri_slot, (uint32_t)g->ntriangles, n, pos, dir, last_hit, out_tri, out_dist,
            nullptr, status, prm);
    }
    KERNEL_CHECK(c);
    return 0;
}

// ---------------------------------------------------------------------------------------
// fused persistent propagate kernel on the wide tree
//
// Replaces chroma/cuda/propagate.cu:217-319 + the host step loop chroma/gpu/photon.py:193-259.
// Every lane owns one photon from its first step to its last; the warp votes between
//   physics phase — when >= `refill` lanes have finished their search (or nothing else can run):
//                   propagate.cu:277-300 for those lanes, then retire / start the next search;
//                   empty lanes are refilled from the queue head (one atomicAdd per warp)
//   search phase  — `iters` search steps (wide.cuh) for the lanes that are searching
// Per photon the sequence of random draws and physics is that of the reference.
// ---------------------------------------------------------------------------------------
#include "wide_prop_kernel.cuh"

#if CHROMA_AB_KERNELS
#include "wide_prop2.cuh"
#endif

int chroma_wide_propagate(chroma_photons *p, chroma_geometry *g, chroma_rng *rng, uint64_t offset,
                          uint64_t n, int max_steps, int use_weights, int scatter_first,
                          uint64_t *nan_aborted) {
    chroma_ctx *c = p->ctx;
    const chroma_ctx::Knobs &kn = c->knobs;
    size_t smem = (size_t)WIDE_SSTACK * WPROP_THREADS * sizeof(uint2) +
                  (size_t)g->view.opt.smem_floats * sizeof(float) +
                  (WIDE_LEAF_SPREAD ? (WPROP_THREADS / 32) * WSPREAD_BYTES_PER_WARP : 0);
    if (!c->wide_attr_set) {   // per context (= per device): the opt-in is a per-device function attribute
        CU_CHECK(cudaFuncSetAttribute(wide_propagate_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU_CHECK(cudaFuncSetAttribute(wide_propagate_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
#if CHROMA_AB_KERNELS
        CU_CHECK(cudaFuncSetAttribute(wide_propagate2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU_CHECK(cudaFuncSetAttribute(wide_propagate2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
#endif
        c->wide_attr_set = true;
    }
    // scratch words: [4..5] queue head (u64), [6] nan count, [7] status
    CU_CHECK(cudaMemsetAsync(c->d_scratch + 4, 0, 16, c->stream));
    WPropParams prm;
    prm.offset = offset; prm.n = n;
    prm.seed = rng->seed;
    prm.id_base = (uint32_t)p->id_base;
    prm.epoch = rng->epoch++;
    prm.max_steps = max_steps; prm.use_weights = use_weights; prm.scatter_first = scatter_first;
    prm.refill = kn.refill;            // r02c sweep: 8 -> 259, 16 -> 310, 20 -> 318 Mphot/s
    prm.iters = kn.iters;
    prm.leaf_thresh = kn.leaf_thresh;
    prm.watchdog_spins = kn.watchdog_spins;
    prm.head = (unsigned long long *)(c->d_scratch + 4);
    prm.nan_count = c->d_scratch + 6;
    prm.status = c->d_scratch + 7;
    // CHROMA_PROP_V=5: two photon slots per lane (wide_prop2.cuh; A/B builds only)
#if CHROMA_AB_KERNELS
    const bool two = kn.prop_v >= 5 && !g->instanced;
#endif
    // physics-heavy geometries (reemitting materials, thin-film / WLS / dichroic surfaces) run the build of
    // the same kernel with the random draws out of line (wide_prop_kernel.cuh); CHROMA_RNG_OOL=0|1 forces it
    const bool ool = kn.rng_ool < 0 ? g->physics_heavy : kn.rng_ool != 0;
    int occ = 0;
    if (g->instanced) {
        int rc = chroma_inst_prop_occupancy(c, ool, &smem, &occ);
        if (rc) return rc;
    } else if (ool) {
        int rc = chroma_wide_ool_occupancy(c, smem, &occ);
        if (rc) return rc;
    }
#if CHROMA_AB_KERNELS
    else if (two) CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, wide_propagate2_kernel<false>, WPROP_THREADS, smem));
#endif
    else CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, wide_propagate_kernel<false>, WPROP_THREADS, smem));
    if (occ < 1) occ = 1;
#if CHROMA_AB_KERNELS
    if (two) {
        size_t need = (size_t)c->sm_count * occ * WPROP_THREADS * 2 * W2_WORDS * sizeof(uint32_t);
        if (c->side_bytes < need) {
            if (c->d_side) CU_CHECK(cudaFree(c->d_side));
            c->d_side = nullptr; c->side_bytes = 0;
            CU_CHECK(cudaMalloc(&c->d_side, need));
            c->side_bytes = need;
        }
    }
#endif
    auto launch = [&](bool staged) {
        unsigned grid = grid_for(prm.n, WPROP_THREADS, c, (unsigned)occ);
        if (g->instanced) {
            chroma_inst_prop_launch(g, p, prm, ool, staged, grid, smem);
        } else if (ool) {
            chroma_wide_ool_launch(g, p, prm, staged, grid, smem);
#if CHROMA_AB_KERNELS
        } else if (two) {
            if (staged) wide_propagate2_kernel<true><<<grid, WPROP_THREADS, smem, c->stream>>>(g->wview, g->wide, p->v, prm, c->d_side);
            else wide_propagate2_kernel<false><<<grid, WPROP_THREADS, smem, c->stream>>>(g->wview, g->wide, p->v, prm, c->d_side);
#endif
        } else {
            if (staged) wide_propagate_kernel<true><<<grid, WPROP_THREADS, smem, c->stream>>>(g->wview, g->wide, p->v, prm);
            else wide_propagate_kernel<false><<<grid, WPROP_THREADS, smem, c->stream>>>(g->wview, g->wide, p->v, prm);
        }
    };
    // Groups are allocated level by level, so the head of the array is the top of the tree: an
    // access-policy window marks it persisting in L2 (CHROMA_L2_WINDOW_MB, default 64, 0 = off; r01x: +1 %).
    // The window belongs to (stream, geometry): it is re-applied whenever the geometry this stream
    // last served differs from `g`, and withdrawn in chroma_geometry_destroy.
    if (c->l2_window_geom != g || g->l2_window_mb != kn.l2_window_mb) {
        cudaStreamAttrValue av{};
        size_t bytes = (size_t)kn.l2_window_mb << 20, all = (size_t)g->wide.ngroups * 128;
        if (bytes > all) bytes = all;
        if (bytes > c->l2_persist_max) bytes = c->l2_persist_max;
        av.accessPolicyWindow.base_ptr = (void *)g->d_wgroups;
        av.accessPolicyWindow.num_bytes = bytes;
        av.accessPolicyWindow.hitRatio = 1.0f;
        av.accessPolicyWindow.hitProp = bytes ? cudaAccessPropertyPersisting : cudaAccessPropertyNormal;
        av.accessPolicyWindow.missProp = bytes ? cudaAccessPropertyStreaming : cudaAccessPropertyNormal;
        cudaStreamSetAttribute(c->stream, cudaStreamAttributeAccessPolicyWindow, &av);
        cudaGetLastError();
        if (c->l2_window_geom && c->l2_window_geom != g) c->l2_window_geom->l2_window_stream = nullptr;
        c->l2_window_geom = g;
        g->l2_window_mb = kn.l2_window_mb;
        g->l2_window_stream = c->stream;
    }
    // a staged upload may still be in flight (engine_photons.cu): if this call covers it, the
    // kernel reads the photons from the stage as the copy engine delivers them
    prm.sv = stage_view_defaults();
    prm.sv_offset = 0;
    prm.ready = nullptr;
    bool direct = false;
    if (p->stage) {
        if (offset <= p->st_offset && offset + n >= p->st_offset + p->st_n && kn.upload_direct) {
            direct = true;
        } else {
            int rc = photons_join(p);
            if (rc) return rc;
        }
    }
    if (direct && (offset < p->st_offset || offset + n > p->st_offset + p->st_n)) {
        // photons of the range outside the stage are already records: run them first, normally
        if (offset < p->st_offset) {
            prm.offset = offset; prm.n = p->st_offset - offset;
            launch(false);
            KERNEL_CHECK(c);
            CU_CHECK(cudaMemsetAsync(c->d_scratch + 4, 0, 8, c->stream));
        }
        if (offset + n > p->st_offset + p->st_n) {
            prm.offset = p->st_offset + p->st_n; prm.n = offset + n - prm.offset;
            launch(false);
            KERNEL_CHECK(c);
            CU_CHECK(cudaMemsetAsync(c->d_scratch + 4, 0, 8, c->stream));
        }
        prm.offset = p->st_offset; prm.n = p->st_n;
    }
    if (direct) {
        prm.sv = p->sv;
        prm.sv_offset = p->st_offset;
        prm.ready = p->d_ready;
    }
    launch(direct);
    KERNEL_CHECK(c);
    if (direct) photons_drop_stage(p);   // stream-ordered: the stage is freed after the kernel
    // status word of this launch: bit 0 search-stack overflow, bit 1 staged upload stalled.  With
    // `nan_aborted` the call synchronises and reports it now; without, the word is folded into the
    // context's sticky status by the next synchronising call (chroma_ctx_synchronize, downloads).
    c->status_pending = true;
    if (nan_aborted) {
        CU_CHECK(cudaMemcpyAsync(c->h_scratch + 4, c->d_scratch + 4, 16, cudaMemcpyDeviceToHost, c->stream));
        CU_CHECK(cudaStreamSynchronize(c->stream));
        *nan_aborted = c->h_scratch[6];
        c->status_pending = false;
        return chroma_status_to_error(c->h_scratch[7]);
    }
    return 0;
}

int chroma_status_to_error(uint32_t status) {
    if (status & 2u) {
        chroma_set_error("photon upload stalled: the propagate kernel gave up waiting for staged photons");
        return CHROMA_ERR_STATE;
    }
    if (status & 1u) {
        chroma_set_error("wide tree search stack overflow (> %d entries)", WIDE_STACK);
        return CHROMA_ERR_OVERFLOW;
    }
    return 0;
}

// ---------------------------------------------------------------------------------------
// chroma_intersect_host_static: chroma_intersect_host for callers that may not allocate.
//
// No counterpart in the reference; added for callers that
// cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
// it).  chroma_intersect_host takes its ray / result temporaries from the stream-ordered pool and
// its queue head + status word from the context's scratch; here ALL of that lives in one file-scope
// device array: nothing is allocated or freed, by this function or by anything it calls.
//
//   layout of g_static_scratch:  [0,8) queue head   [8,12) status word   [256, ...) ray staging:
//     pos[MAX][3] f32 | dir[MAX][3] f32 | last_hit[MAX] i32 | triangle[MAX] i32 | distance[MAX] f32
//
// Every GPU that loads the library holds the array (256 + 36 B x CHROMA_STATIC_MAX_RAYS = 9.0 MiB)
// until the process exits.  All calls on a device share it and it keeps the last call's contents:
// head and status are cleared by every call, the staging regions are overwritten before they are
// read.  Two calls that overlap on the GPU (two contexts / streams on one device) race.
// ---------------------------------------------------------------------------------------
#define STATIC_RAY_BYTES 36u
static __device__ __align__(256) unsigned char g_static_scratch[256 + (size_t)CHROMA_STATIC_MAX_RAYS * STATIC_RAY_BYTES];

// In host code `g_static_scratch` is only a stand-in for the symbol: the device address is asked for
// once per device (with that device current) and kept here.
static unsigned char *g_static_addr[64];
static std::mutex g_static_mutex;

static int static_scratch_address(chroma_ctx *c, unsigned char **out) {
    if (c->device < 0 || c->device >= 64) {
        chroma_set_error("chroma_intersect_host_static: device ordinal %d out of range", c->device);
        return CHROMA_ERR_ARG;
    }
    std::lock_guard<std::mutex> lk(g_static_mutex);
    if (!g_static_addr[c->device]) {
        void *ptr = nullptr;
        CU_CHECK(cudaGetSymbolAddress(&ptr, g_static_scratch));   // ctx_bind made c->device current
        g_static_addr[c->device] = (unsigned char *)ptr;
    }
    *out = g_static_addr[c->device];
    return 0;
}

extern "C" int chroma_intersect_host_static(chroma_geometry *g, uint64_t n, const float *pos, const float *dir,
                                            const int32_t *last_hit, int32_t *out_triangle,
                                            float *out_distance) {
    ARG_CHECK(g && (n == 0 || (pos && dir && out_triangle && out_distance)));
    chroma_ctx *c = g->ctx;
    CTX_BIND(c);
    if (n > (uint64_t)CHROMA_STATIC_MAX_RAYS) {
        chroma_set_error("chroma_intersect_host_static: %llu rays exceed the static scratch (at most %u rays per call)",
                         (unsigned long long)n, (unsigned)CHROMA_STATIC_MAX_RAYS);
        return CHROMA_ERR_ARG;
    }
    if (g->instanced || !g->wide.groups) {
        chroma_set_error("chroma_intersect_host_static: flat geometries with the engine's wide tree only");
        return CHROMA_ERR_STATE;
    }
    if (n == 0) return 0;
    unsigned char *base = nullptr;
    int rc = static_scratch_address(c, &base);
    if (rc) return rc;
    const size_t cap = CHROMA_STATIC_MAX_RAYS;
    float *d_pos = (float *)(base + 256), *d_dir = d_pos + 3 * cap;
    int32_t *d_last = (int32_t *)(d_dir + 3 * cap), *d_tri = d_last + cap;
    float *d_dist = (float *)(d_tri + cap);
    uint32_t *status = (uint32_t *)(base + 8);
    WIsectParams prm;
    prm.head = (unsigned long long *)base;
    CU_CHECK(cudaMemsetAsync(base, 0, 16, c->stream));            // head + status start at zero
    CU_CHECK(cudaMemcpyAsync(d_pos, pos, n * 12, cudaMemcpyHostToDevice, c->stream));
    CU_CHECK(cudaMemcpyAsync(d_dir, dir, n * 12, cudaMemcpyHostToDevice, c->stream));
    if (last_hit) CU_CHECK(cudaMemcpyAsync(d_last, last_hit, n * 4, cudaMemcpyHostToDevice, c->stream));
    prm.refill = c->knobs.isect_refill;
    prm.iters = c->knobs.isect_iters;
    prm.leaf_thresh = c->knobs.leaf_thresh;
    size_t smem = (size_t)WIDE_SSTACK * WISECT_THREADS * sizeof(uint2) + (WIDE_LEAF_SPREAD ? (WISECT_THREADS / 32) * WSPREAD_BYTES_PER_WARP : 0);
    int occ = 0;
    CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, wide_intersect_kernel<false>, WISECT_THREADS, smem));
    if (occ < 1) occ = 1;
    unsigned grid = grid_for(n, WISECT_THREADS, c, (unsigned)occ);
    wide_intersect_kernel<false><<<grid, WISECT_THREADS, smem, c->stream>>>(
        g->wide, g->d_tri_slot, (uint32_t)g->ntriangles, n, d_pos, d_dir, last_hit ? d_last : nullptr, d_tri, d_dist,
        nullptr, status, prm);
    KERNEL_CHECK(c);
    CU_CHECK(cudaMemcpyAsync(out_triangle, d_tri, n * 4, cudaMemcpyDeviceToHost, c->stream));
    CU_CHECK(cudaMemcpyAsync(out_distance, d_dist, n * 4, cudaMemcpyDeviceToHost, c->stream));
    CU_CHECK(cudaMemcpyAsync(c->h_scratch + 9, status, 4, cudaMemcpyDeviceToHost, c->stream));
    CU_CHECK(cudaStreamSynchronize(c->stream));
    if (c->h_scratch[9] & 1u) {
        chroma_set_error("wide tree search stack overflow (> %d entries)", WIDE_STACK);
        return CHROMA_ERR_OVERFLOW;
    }
    return 0;
}
