// engine_photons.cu — intersect / propagate kernels and the GPUPhotons C-ABI.
//
// Replaces chroma/cuda/mesh.h:123-151 (distance_to_mesh), chroma/cuda/propagate.cu (all
// kernels) and the host loop of chroma/gpu/photon.py:193-259.
//
// B200 design (vs the reference's one-thread-per-photon, one-launch-per-step scheme):
//   * photon state is four 16-byte records per photon (pos|t, dir|wavelength, pol|weight,
//     flags|last_hit|evidx|channel): every load/store is a coalesced 128-bit access;
//   * ONE persistent launch runs the whole step loop.  Warps pull photons from a global
//     queue head; when >= `refill` lanes of a warp have retired their photon the warp
//     ballots the free lanes and refills them with one atomicAdd (warp-ballot compaction of
//     the live set without a host round trip or a state round trip through HBM);
//   * the top of the BVH and the wavelength tables are staged in shared memory, the rest of
//     the node array sits behind an L2 access-policy window (engine_core.cu);
//   * RNG is counter-based (Philox4x32-10 keyed by photon id and call epoch): no state
//     traffic, results independent of launch shape and of how photons are sharded.
#include <stdlib.h>
#include <algorithm>
#include <thread>
#include <vector>
#include "physics.cuh"

#define PROP_THREADS 256
#define TRY_RC(x)                                                                          \
    do {                                                                                   \
        int _r = (x);                                                                      \
        if (_r) return _r;                                                                 \
    } while (0)

#define ISECT_THREADS 256

static uint32_t env_u32(const char *name, uint32_t dflt) {
    const char *s = getenv(name);
    if (!s || !*s) return dflt;
    return (uint32_t)strtoul(s, nullptr, 10);
}

// cooperative staging of the BVH top and the optics tables into shared memory
__device__ __forceinline__ void stage_smem(const GeomView &g, uint4 *s_nodes, float *s_opt) {
    for (uint32_t i = threadIdx.x; i < g.top_nodes; i += blockDim.x) s_nodes[i] = __ldg(g.nodes + i);
    for (uint32_t i = threadIdx.x; i < g.opt.smem_floats; i += blockDim.x) s_opt[i] = __ldg(g.opt.blob + i);
    __syncthreads();
}


// ---------------------------------------------------------------------------------------
// first-hit intersection (mesh.h:123-151 + the index the reference drops)
// ---------------------------------------------------------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(ISECT_THREADS)
intersect_kernel(GeomView g, uint64_t n, const float *__restrict__ pos,
                 const float *__restrict__ dir, const int32_t *__restrict__ last_hit,
                 int32_t *__restrict__ out_tri, float *__restrict__ out_dist,
                 uint32_t *__restrict__ out_counts, uint32_t *status) {
    extern __shared__ float4 smem_raw[];
    uint4 *s_nodes = (uint4 *)smem_raw;
    float *s_opt = (float *)(s_nodes + g.top_nodes);
    {
        for (uint32_t i = threadIdx.x; i < g.top_nodes; i += blockDim.x) s_nodes[i] = __ldg(g.nodes + i);
        __syncthreads();
    }
    (void)s_opt;
    unsigned stack[CHROMA_STACK_WORDS];
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t id = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; id < n; id += stride) {
        float3 o = f3(pos[3 * id], pos[3 * id + 1], pos[3 * id + 2]);
        float3 d = f3(dir[3 * id], dir[3 * id + 1], dir[3 * id + 2]);
        // mesh.h:141  direction /= norm(direction)  (fast-math: sqrt.approx, div.approx)
        float nrm = sqrtf(ref_dot(d, d));
        d = f3(__fdividef(d.x, nrm), __fdividef(d.y, nrm), __fdividef(d.z, nrm));
        int last_slot = -1;
        if (last_hit) {
            int lt = last_hit[id];
            if (lt >= 0 && (uint32_t)lt < g.ntriangles) last_slot = g.tri_slot[lt];
        }
        RayCtx r;
        ray_setup(r, o, d);
        float dist;
        uint32_t nc = 0, tc = 0;
        bool ovf = false;
        int slot = intersect_mesh_slot<COUNT>(g, s_nodes, r, last_slot, dist, stack, &nc, &tc, ovf);
        int tri = -1;
        if (slot >= 0) tri = (int)__float_as_uint(__ldg(g.tris + (size_t)slot * 3 + 2).z);
        out_tri[id] = tri;
        out_dist[id] = dist;
        if (COUNT) {
            out_counts[2 * id] = nc;
            out_counts[2 * id + 1] = tc;
        }
        if (ovf) atomicOr(status, 1u);
    }
}


#if CHROMA_AB_KERNELS
// v2: persistent, warp-scheduled (see traverse.cuh "Warp-scheduled form"): lanes pull rays from
// a global head, node tests and triangle tests run as separate warp-wide phases
struct IsectParams {
    unsigned long long *head;
    int refill, leaf_thresh, node_iters;
};

__global__ void __launch_bounds__(ISECT_THREADS)
intersect_kernel_v2(GeomView g, uint64_t n, const float *__restrict__ pos,
                    const float *__restrict__ dir, const int32_t *__restrict__ last_hit,
                    int32_t *__restrict__ out_tri, float *__restrict__ out_dist, uint32_t *status,
                    IsectParams prm) {
    extern __shared__ float4 smem_raw[];
    uint4 *s_nodes = (uint4 *)smem_raw;
    for (uint32_t i = threadIdx.x; i < g.top_nodes; i += blockDim.x) s_nodes[i] = __ldg(g.nodes + i);
    __syncthreads();
    unsigned stack[CHROMA_STACK_WORDS];
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    bool have = false, exhausted = false, overflow = false;
    float3 o = f3(0, 0, 0), d = f3(0, 0, 1);
    uint64_t id = 0;
    uint32_t dummy = 0;
    Trav tv;
    tv.leaf = -1; tv.cur = tv.end = 0; tv.sp = -1;
    while (true) {
        const bool nodew = have && trav_node_work(tv);
        const bool leafw = have && tv.leaf >= 0;
        const bool resolve = have && !nodew && !leafw;
        const unsigned mN = __ballot_sync(0xffffffffu, nodew);
        const unsigned mL = __ballot_sync(0xffffffffu, leafw);
        const unsigned mR = __ballot_sync(0xffffffffu, resolve);
        const unsigned mH = __ballot_sync(0xffffffffu, have);
        if (mH == 0u && exhausted) break;
        const int nidle = __popc(mR) + (exhausted ? 0 : 32 - __popc(mH));
        if ((mN | mL) == 0u || nidle >= prm.refill) {
            if (resolve) {
                int tri = -1;
                if (tv.hit >= 0) tri = (int)__float_as_uint(__ldg(g.tris + (size_t)tv.hit * 3 + 2).z);
                out_tri[id] = tri;
                out_dist[id] = tv.hit >= 0 ? tv.dist : -1.0f;
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
                                if (lt >= 0 && (uint32_t)lt < g.ntriangles) last_slot = g.tri_slot[lt];
                            }
                            trav_begin(tv, g, s_nodes, o, d, last_slot);
                            have = true;
                        }
                    }
                }
            }
        } else if (mN == 0u || __popc(mL) >= prm.leaf_thresh) {
            if (leafw) trav_leaf_step(tv, g, o, d);
        } else {
            for (int it = 0; it < prm.node_iters; it++)
                if (have && trav_node_work(tv)) trav_node_step<false>(tv, g, s_nodes, o, d, stack, overflow, dummy);
        }
    }
    if (overflow) atomicOr(status, 1u);
}

#endif  // CHROMA_AB_KERNELS

static int check_overflow(chroma_ctx *c, uint32_t *d_status) {
    CU_CHECK(cudaMemcpyAsync(c->h_scratch + 8, d_status, 4, cudaMemcpyDeviceToHost, c->stream));
    CU_CHECK(cudaStreamSynchronize(c->stream));
    if (c->h_scratch[8] & 1u) {
        chroma_set_error("BVH traversal stack overflow (> %d pending sibling groups)", CHROMA_STACK_WORDS);
        return CHROMA_ERR_OVERFLOW;
    }
    return 0;
}

template <bool COUNT>
static int launch_intersect(chroma_geometry *g, uint64_t n, const float *pos, const float *dir,
                            const int32_t *last_hit, int32_t *out_tri, float *out_dist,
                            uint32_t *out_counts) {
    chroma_ctx *c = g->ctx;
    if (n == 0) return 0;
    size_t smem = (size_t)g->view.top_nodes * sizeof(uint4);
    static bool attr_set[2] = {false, false};
    if (!attr_set[COUNT]) {
        CU_CHECK(cudaFuncSetAttribute(intersect_kernel<COUNT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        attr_set[COUNT] = true;
    }
    // default: near-first search of the engine's wide tree (engine_wide.cu) or of the two-level tree;
    // the counting entry walks the reference-format tree (its visit counts define the algorithmic
    // bytes) unless CHROMA_COUNT_WIDE=1 asks for the engine tree's own (groups, triangle tests).
    // CHROMA_ISECT_V=1: one ray per lane on the reference-format tree in the reference's order (debug).
    const chroma_ctx::Knobs &kn = c->knobs;
    if (g->instanced || (COUNT ? kn.count_wide == 1 : kn.isect_v != 1 && (kn.isect_v >= 3 || !CHROMA_AB_KERNELS)))
        return chroma_wide_intersect(g, n, pos, dir, last_hit, out_tri, out_dist, out_counts);
    uint32_t *status = c->d_scratch + 2;
    CU_CHECK(cudaMemsetAsync(status, 0, 4, c->stream));
#if CHROMA_AB_KERNELS
    if (!COUNT && kn.isect_v != 1) {
        static bool attr2 = false;
        if (!attr2) {
            CU_CHECK(cudaFuncSetAttribute(intersect_kernel_v2, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
            attr2 = true;
        }
        IsectParams prm;
        prm.head = (unsigned long long *)(c->d_scratch + 10);
        CU_CHECK(cudaMemsetAsync(prm.head, 0, 8, c->stream));
        prm.refill = (int)env_u32("CHROMA_ISECT_REFILL", 8);
        prm.leaf_thresh = (int)env_u32("CHROMA_LEAF_THRESH", 4);
        prm.node_iters = (int)env_u32("CHROMA_NODE_ITERS", 8);
        if (prm.refill < 1) prm.refill = 1;
        if (prm.leaf_thresh < 1) prm.leaf_thresh = 1;
        if (prm.node_iters < 1) prm.node_iters = 1;
        int occ = 0;
        CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, intersect_kernel_v2, ISECT_THREADS, smem));
        if (occ < 1) occ = 1;
        unsigned grid = grid_for(n, ISECT_THREADS, c, (unsigned)occ);
        intersect_kernel_v2<<<grid, ISECT_THREADS, smem, c->stream>>>(g->view, n, pos, dir, last_hit, out_tri,
                                                                       out_dist, status, prm);
        KERNEL_CHECK(c);
        return 0;
    }
#endif
    unsigned per_sm = 6;
    unsigned grid = grid_for(n, ISECT_THREADS, c, per_sm);
    intersect_kernel<COUNT><<<grid, ISECT_THREADS, smem, c->stream>>>(
        g->view, n, pos, dir, last_hit, out_tri, out_dist, out_counts, status);
    KERNEL_CHECK(c);
    return 0;
}

extern "C" int chroma_intersect_device(chroma_geometry *g, uint64_t n, const float *pos,
                                       const float *dir, const int32_t *last_hit,
                                       int32_t *out_triangle, float *out_distance) {
    ARG_CHECK(g && (n == 0 || (pos && dir && out_triangle && out_distance)));
    CTX_BIND(g->ctx);
    return launch_intersect<false>(g, n, pos, dir, last_hit, out_triangle, out_distance, nullptr);
}

extern "C" int chroma_intersect_count_device(chroma_geometry *g, uint64_t n, const float *pos,
                                             const float *dir, const int32_t *last_hit,
                                             int32_t *out_triangle, float *out_distance,
                                             uint32_t *out_counts) {
    ARG_CHECK(g && (n == 0 || (pos && dir && out_triangle && out_distance && out_counts)));
    CTX_BIND(g->ctx);
    return launch_intersect<true>(g, n, pos, dir, last_hit, out_triangle, out_distance, out_counts);
}

extern "C" int chroma_intersect_host(chroma_geometry *g, uint64_t n, const float *pos,
                                     const float *dir, const int32_t *last_hit,
                                     int32_t *out_triangle, float *out_distance) {
    ARG_CHECK(g && (n == 0 || (pos && dir && out_triangle && out_distance)));
    chroma_ctx *c = g->ctx;
    CTX_BIND(c);
    if (n == 0) return 0;
    float *d_pos = nullptr, *d_dir = nullptr, *d_dist = nullptr;
    int32_t *d_last = nullptr, *d_tri = nullptr;
    int rc = 0;
    do {
        // per-call temporaries come from the stream-ordered pool (see chroma_malloc)
        if (cudaMallocAsync(&d_pos, n * 12, c->stream) || cudaMallocAsync(&d_dir, n * 12, c->stream) ||
            cudaMallocAsync(&d_dist, n * 4, c->stream) || cudaMallocAsync(&d_tri, n * 4, c->stream) ||
            (last_hit && cudaMallocAsync(&d_last, n * 4, c->stream))) {
            chroma_set_error("cudaMalloc: %s", cudaGetErrorString(cudaGetLastError()));
            rc = CHROMA_ERR_CUDA; break;
        }
        cudaMemcpyAsync(d_pos, pos, n * 12, cudaMemcpyHostToDevice, c->stream);
        cudaMemcpyAsync(d_dir, dir, n * 12, cudaMemcpyHostToDevice, c->stream);
        if (last_hit) cudaMemcpyAsync(d_last, last_hit, n * 4, cudaMemcpyHostToDevice, c->stream);
        if ((rc = launch_intersect<false>(g, n, d_pos, d_dir, d_last, d_tri, d_dist, nullptr))) break;
        cudaMemcpyAsync(out_triangle, d_tri, n * 4, cudaMemcpyDeviceToHost, c->stream);
        cudaMemcpyAsync(out_distance, d_dist, n * 4, cudaMemcpyDeviceToHost, c->stream);
        if ((rc = check_overflow(c, c->d_scratch + 2))) break;
    } while (0);
    if (d_pos) cudaFreeAsync(d_pos, c->stream);
    if (d_dir) cudaFreeAsync(d_dir, c->stream);
    if (d_dist) cudaFreeAsync(d_dist, c->stream);
    if (d_tri) cudaFreeAsync(d_tri, c->stream);
    if (d_last) cudaFreeAsync(d_last, c->stream);
    return rc;
}

// ---------------------------------------------------------------------------------------
// RNG handle
// ---------------------------------------------------------------------------------------
extern "C" int chroma_rng_create(chroma_ctx *c, uint64_t size, uint64_t seed, chroma_rng **out) {
    ARG_CHECK(c && out);
    (void)size;
    chroma_rng *r = new chroma_rng();
    r->ctx = c;
    r->seed = seed;
    r->epoch = 0;
    *out = r;
    return 0;
}
extern "C" int chroma_rng_destroy(chroma_rng *r) {
    delete r;
    return 0;
}
extern "C" int chroma_rng_get_epoch(chroma_rng *r, uint32_t *epoch) {
    ARG_CHECK(r && epoch);
    *epoch = r->epoch;
    return 0;
}
extern "C" int chroma_rng_set_epoch(chroma_rng *r, uint32_t epoch) {
    ARG_CHECK(r);
    r->epoch = epoch;
    return 0;
}

__global__ void fill_uniform_kernel(uint64_t n, float *out, uint64_t seed, uint32_t epoch) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const RngKey K = rng_key(seed, epoch, 2u);
        Rng g;
        rng_init(g, (uint32_t)i);
        out[i] = rng_uniform(g, K);
    }
}

extern "C" int chroma_rng_fill_uniform(chroma_rng *r, uint64_t n, float *out_device) {
    ARG_CHECK(r && (n == 0 || out_device));
    chroma_ctx *c = r->ctx;
    CTX_BIND(c);
    if (n == 0) return 0;
    fill_uniform_kernel<<<grid_for(n, 256, c, 8), 256, 0, c->stream>>>(n, out_device, r->seed, r->epoch++);
    KERNEL_CHECK(c);
    return 0;
}

// ---------------------------------------------------------------------------------------
// photon containers
// ---------------------------------------------------------------------------------------
extern "C" int chroma_photons_create(chroma_ctx *c, uint64_t capacity, uint64_t id_base,
                                     chroma_photons **out) {
    CTX_BIND(c);
    ARG_CHECK(out != nullptr);
    chroma_photons *p = new chroma_photons();
    p->ctx = c;
    p->capacity = capacity;
    p->id_base = id_base;
    size_t nb = (capacity ? capacity : 1) * 16;
    // stream-ordered pool allocation: a container per batch costs microseconds, not a cudaMalloc
    if (cudaMallocAsync(&p->v.pt, nb, c->stream) || cudaMallocAsync(&p->v.dw, nb, c->stream) ||
        cudaMallocAsync(&p->v.pw, nb, c->stream) || cudaMallocAsync(&p->v.st, nb, c->stream)) {
        chroma_set_error("cudaMalloc photons (%llu): %s", (unsigned long long)capacity,
                         cudaGetErrorString(cudaGetLastError()));
        chroma_photons_destroy(p);
        return CHROMA_ERR_CUDA;
    }
    *out = p;
    return 0;
}

extern "C" int chroma_photons_destroy(chroma_photons *p) {
    if (!p) return 0;
    cudaSetDevice(p->ctx->device);
    photons_join(p);
    cudaStreamSynchronize(p->ctx->stream);
    if (p->v.pt) cudaFreeAsync(p->v.pt, p->ctx->stream);
    if (p->v.dw) cudaFreeAsync(p->v.dw, p->ctx->stream);
    if (p->v.pw) cudaFreeAsync(p->v.pw, p->ctx->stream);
    if (p->v.st) cudaFreeAsync(p->v.st, p->ctx->stream);
    if (p->wave.aux) cudaFree(p->wave.aux);
    if (p->wave.hit) cudaFree(p->wave.hit);
    if (p->wave.queue[0]) cudaFree(p->wave.queue[0]);
    if (p->wave.queue[1]) cudaFree(p->wave.queue[1]);
    if (p->wave.ctr) cudaFree(p->wave.ctr);
    delete p;
    return 0;
}

extern "C" int chroma_photons_size(chroma_photons *p, uint64_t *capacity) {
    ARG_CHECK(p && capacity);
    *capacity = p->capacity;
    return 0;
}

// host AoS (event.Photons) -> device records.  `stage` holds the raw host arrays on device.
struct StageOut {
    float *pos, *dir, *pol, *wl, *t, *weights;
    int32_t *last_hit, *channel;
    uint32_t *flags, *evidx;
};

__global__ void pack_photons_kernel(PhotonView v, uint64_t offset, uint64_t n, StageView s) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        uint64_t j = offset + i;
        v.pt[j] = s.pos ? make_float4(s.pos[3 * i], s.pos[3 * i + 1], s.pos[3 * i + 2], s.t ? s.t[i] : s.t_v)
                        : make_float4(s.pos_v[0], s.pos_v[1], s.pos_v[2], s.t ? s.t[i] : s.t_v);
        v.dw[j] = make_float4(s.dir[3 * i], s.dir[3 * i + 1], s.dir[3 * i + 2], s.wl[i]);
        v.pw[j] = make_float4(s.pol[3 * i], s.pol[3 * i + 1], s.pol[3 * i + 2],
                              s.weights ? s.weights[i] : s.weight_v);
        v.st[j] = make_uint4(s.flags ? s.flags[i] : s.flags_v, (uint32_t)(s.last_hit ? s.last_hit[i] : s.last_hit_v),
                             s.evidx ? s.evidx[i] : s.evidx_v, 0xFFFFFFFFu);
    }
}

__global__ void unpack_photons_kernel(PhotonView v, uint64_t offset, uint64_t n, StageOut s) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        uint64_t j = offset + i;
        float4 a = v.pt[j], b = v.dw[j], c = v.pw[j];
        uint4 d = v.st[j];
        if (s.pos) { s.pos[3 * i] = a.x; s.pos[3 * i + 1] = a.y; s.pos[3 * i + 2] = a.z; }
        if (s.t) s.t[i] = a.w;
        if (s.dir) { s.dir[3 * i] = b.x; s.dir[3 * i + 1] = b.y; s.dir[3 * i + 2] = b.z; }
        if (s.wl) s.wl[i] = b.w;
        if (s.pol) { s.pol[3 * i] = c.x; s.pol[3 * i + 1] = c.y; s.pol[3 * i + 2] = c.z; }
        if (s.weights) s.weights[i] = c.w;
        if (s.flags) s.flags[i] = d.x;
        if (s.last_hit) s.last_hit[i] = (int32_t)d.y;
        if (s.evidx) s.evidx[i] = d.z;
        if (s.channel) s.channel[i] = (int32_t)d.w;
    }
}

// Parallel host copy of one chunk of every pageable field into a slot of the page-locked ring: one set
// of threads per chunk, each copying its share of the chunk's photons of every field.
struct HostCopy { char *dst; const char *src; size_t elem; };
static void host_copy_chunk(const HostCopy *f, int nfields, size_t nphotons, int nthreads) {
    auto part = [=](int k, int parts) {
        const size_t per = ((nphotons + parts - 1) / parts + 15) & ~(size_t)15, b = per * k;
        if (b >= nphotons) return;
        const size_t m = std::min(per, nphotons - b);
        for (int q = 0; q < nfields; q++) memcpy(f[q].dst + b * f[q].elem, f[q].src + b * f[q].elem, m * f[q].elem);
    };
    if (nphotons < (1u << 16) || nthreads <= 1) { part(0, 1); return; }
    std::vector<std::thread> th;
    for (int k = 1; k < nthreads; k++) th.emplace_back(part, k, nthreads);
    part(0, nthreads);
    for (auto &t : th) t.join();
}

static bool host_pointer_is_pinned(const void *ptr) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, ptr) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeHost;
}

extern "C" int chroma_photons_upload_arrays(chroma_photons *p, uint64_t offset, uint64_t n,
                                            const chroma_photon_arrays *a) {
    ARG_CHECK(p && a);
    chroma_ctx *c = p->ctx;
    CTX_BIND(c);
    ARG_CHECK(offset + n <= p->capacity);
    if (n == 0) return 0;
    ARG_CHECK(a->dir && a->pol && a->wavelengths);
    TRY_RC(photons_join(p));
    TRY_RC(uploader_join(c));      // one staged upload per context at a time (the ring and its events are shared)
    // The host arrays go to a device stage as they are (<= 60 B/photon, the event.Photons payload;
    // fields given by value are not sent at all), chunk by chunk on the context's copy stream; after
    // each chunk the copy engine also writes the count of photons landed so far to a device word.
    // chroma_photons_propagate polls that word and reads photons straight from the stage, so the
    // search runs while the rest is still crossing PCIe and no SM is needed to receive data (a packing
    // kernel could not be scheduled next to the persistent search kernel).  Any other consumer first
    // calls photons_join(), which waits and packs the 16-byte records.
    // Page-locked sources are read by DMA directly.  Pageable sources go through the context's
    // page-locked ring: host threads copy chunk k+1 into a free slot while the DMA of chunk k runs
    // (cudaMemcpyAsync from pageable memory would be staged by the runtime, one call at a time, at a
    // fifth of the PCIe rate: DESIGN.md section 3).
    uint64_t k = (uint64_t)c->knobs.upload_chunks;
    const uint64_t min_chunk = 1u << 17;
    if (k < 1) k = 1;
    if (n / k < min_chunk) k = n / min_chunk ? n / min_chunk : 1;
    if (k > READY_RING / 4) k = READY_RING / 4;     // ready-ring slots of one upload are never reused while it runs
    const uint64_t per = (n + k - 1) / k;
    struct Field { const void *src; size_t elem; const void **view; bool pinned; };
    StageView sv = stage_view_defaults();
    for (int q = 0; q < 3; q++) sv.pos_v[q] = a->pos_value[q];
    sv.t_v = a->t_value; sv.weight_v = a->weight_value; sv.flags_v = a->flags_value;
    sv.evidx_v = a->evidx_value; sv.last_hit_v = a->last_hit_value;
    Field fields[9] = {{a->pos, 12, (const void **)&sv.pos, true}, {a->dir, 12, (const void **)&sv.dir, true},
                       {a->pol, 12, (const void **)&sv.pol, true}, {a->wavelengths, 4, (const void **)&sv.wl, true},
                       {a->t, 4, (const void **)&sv.t, true}, {a->weights, 4, (const void **)&sv.weights, true},
                       {a->last_hit, 4, (const void **)&sv.last_hit, true}, {a->flags, 4, (const void **)&sv.flags, true},
                       {a->evidx, 4, (const void **)&sv.evidx, true}};
    size_t bytes_per_photon = 0, pageable_bytes_per_photon = 0;
    for (Field &f : fields)
        if (f.src) {
            bytes_per_photon += f.elem;
            f.pinned = host_pointer_is_pinned(f.src);      // page-locked fields go by DMA directly,
            if (!f.pinned) pageable_bytes_per_photon += f.elem;   // pageable ones through the ring
        }
    const bool pinned = pageable_bytes_per_photon == 0;
    CU_CHECK(cudaMallocAsync(&p->stage, n * bytes_per_photon, c->stream));
    CU_CHECK(cudaMallocAsync(&p->d_ready, 8, c->stream));
    CU_CHECK(cudaMemsetAsync(p->d_ready, 0, 8, c->stream));
    CU_CHECK(cudaEventRecord(c->copy_fork, c->stream));          // allocations are ordered on c->stream
    CU_CHECK(cudaStreamWaitEvent(c->copy_stream, c->copy_fork, 0));
    {
        size_t off = 0;
        for (const Field &f : fields) {
            *f.view = f.src ? (const void *)(p->stage + off) : nullptr;
            if (f.src) off += n * f.elem;
        }
    }
    p->sv = sv;
    p->st_offset = offset;
    p->st_n = n;
    if (!pinned) {
        const size_t need = per * pageable_bytes_per_photon;
        if (c->stage_slot_bytes < need) {
            CU_CHECK(cudaStreamSynchronize(c->copy_stream));
            if (c->h_stage_ring) CU_CHECK(cudaFreeHost(c->h_stage_ring));
            c->h_stage_ring = nullptr; c->stage_slot_bytes = 0;
            CU_CHECK(cudaMallocHost(&c->h_stage_ring, need * 4));
            c->stage_slot_bytes = need;
            for (cudaEvent_t &e : c->stage_slot_done)
                if (!e) CU_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        }
    }
    // what the chunk loop needs, by value: it may run on a worker thread after this call has returned
    struct Job { const char *src; char *dst; size_t elem; bool pinned; };
    Job jobs[9];
    int njobs = 0;
    for (const Field &f : fields)
        if (f.src) jobs[njobs++] = Job{(const char *)f.src, (char *)*f.view, f.elem, f.pinned};
    CU_CHECK(cudaEventCreateWithFlags(&p->st_done, cudaEventDisableTiming));
    const int nthreads = c->knobs.upload_threads;
    cudaEvent_t st_done = p->st_done;
    unsigned long long *d_ready = p->d_ready;
    auto run_chunks = [=]() -> int {
        for (uint64_t b = 0; b < n; b += per) {
            const uint64_t m = (n - b < per) ? n - b : per;
            char *slot = nullptr;
            uint32_t slot_id = 0;
            if (!pinned) {
                slot_id = c->stage_slot_pos++ & 3u;
                slot = c->h_stage_ring + (size_t)slot_id * c->stage_slot_bytes;
                CU_CHECK(cudaEventSynchronize(c->stage_slot_done[slot_id]));   // its last DMA has finished
                HostCopy hc[9];
                int nf = 0;
                size_t soff = 0;
                for (int q = 0; q < njobs; q++) {
                    if (jobs[q].pinned) continue;
                    hc[nf++] = HostCopy{slot + soff, jobs[q].src + b * jobs[q].elem, jobs[q].elem};
                    soff += m * jobs[q].elem;
                }
                host_copy_chunk(hc, nf, m, nthreads);
            }
            size_t soff = 0;
            for (int q = 0; q < njobs; q++) {
                const Job &f = jobs[q];
                const char *src = f.pinned ? f.src + b * f.elem : slot + soff;
                if (!f.pinned) soff += m * f.elem;
                CU_CHECK(cudaMemcpyAsync(f.dst + b * f.elem, src, m * f.elem, cudaMemcpyHostToDevice, c->copy_stream));
            }
            if (!pinned) CU_CHECK(cudaEventRecord(c->stage_slot_done[slot_id], c->copy_stream));
            unsigned long long *ready = c->h_ready_ring + (c->ready_ring_pos++ % READY_RING);
            *ready = b + m;
            CU_CHECK(cudaMemcpyAsync(d_ready, ready, 8, cudaMemcpyHostToDevice, c->copy_stream));
        }
        CU_CHECK(cudaEventRecord(st_done, c->copy_stream));
        return 0;
    };
    if (a->borrow_sources && !pinned) {
        // The caller keeps the pageable arrays untouched until the next synchronising call (as it must for
        // page-locked ones): the ring is filled by a worker thread, so the caller can launch the propagate
        // kernel at once and the search runs while the host is still copying (r2l: the host copy of a 600 MB
        // bomb takes 17 ms, during which the kernel could not even be launched).  The worker is joined by
        // whoever needs the stage complete (uploader_join).
        chroma_ctx *cc = c;
        const int dev = c->device;
        c->uploader_rc = 0;
        c->uploader = new std::thread([cc, dev, run_chunks]() {
            cudaSetDevice(dev);
            const int rc = run_chunks();
            if (rc) {
                cc->uploader_rc = rc;
                snprintf(cc->uploader_err, sizeof(cc->uploader_err), "%s", chroma_last_error());
            }
        });
        return 0;
    }
    TRY_RC(run_chunks());
    CU_CHECK(cudaGetLastError());
    return 0;
}

extern "C" int chroma_photons_upload(chroma_photons *p, uint64_t offset, uint64_t n,
                                     const float *pos, const float *dir, const float *pol,
                                     const float *wavelengths, const float *t,
                                     const int32_t *last_hit, const uint32_t *flags,
                                     const float *weights, const uint32_t *evidx) {
    ARG_CHECK(n == 0 || (pos && dir && pol && wavelengths && t));
    chroma_photon_arrays a{};
    a.pos = pos; a.dir = dir; a.pol = pol; a.wavelengths = wavelengths; a.t = t; a.weights = weights;
    a.last_hit = last_hit; a.flags = flags; a.evidx = evidx;
    a.weight_value = 1.0f; a.last_hit_value = -1;
    return chroma_photons_upload_arrays(p, offset, n, &a);
}

// wait (on the host) for the worker thread of a borrowed-source upload: after it every copy of the
// upload has been enqueued and its completion event recorded
int uploader_join(chroma_ctx *c) {
    if (!c->uploader) return 0;
    std::thread *t = (std::thread *)c->uploader;
    t->join();
    delete t;
    c->uploader = nullptr;
    if (c->uploader_rc) {
        chroma_set_error("staged upload (worker thread): %s", c->uploader_err);
        const int rc = c->uploader_rc;
        c->uploader_rc = 0;
        return rc;
    }
    return 0;
}

void photons_drop_stage(chroma_photons *p) {
    if (!p->stage) return;
    chroma_ctx *c = p->ctx;
    uploader_join(c);      // st_done must have been recorded before anything waits on it
    cudaStreamWaitEvent(c->stream, p->st_done, 0);
    cudaEventDestroy(p->st_done);
    cudaFreeAsync(p->stage, c->stream);
    cudaFreeAsync(p->d_ready, c->stream);
    p->stage = nullptr; p->d_ready = nullptr; p->st_done = nullptr;
    p->sv = stage_view_defaults();
    p->st_n = 0;
}

int photons_join(chroma_photons *p) {
    if (!p->stage) return 0;
    chroma_ctx *c = p->ctx;
    TRY_RC(uploader_join(c));
    CU_CHECK(cudaStreamWaitEvent(c->stream, p->st_done, 0));
    pack_photons_kernel<<<grid_for(p->st_n, 256, c, 8), 256, 0, c->stream>>>(p->v, p->st_offset, p->st_n, p->sv);
    KERNEL_CHECK(c);
    photons_drop_stage(p);
    return 0;
}

extern "C" int chroma_photons_download(chroma_photons *p, uint64_t offset, uint64_t n, float *pos,
                                       float *dir, float *pol, float *wavelengths, float *t,
                                       int32_t *last_hit, uint32_t *flags, float *weights,
                                       uint32_t *evidx, int32_t *channel) {
    ARG_CHECK(p);
    chroma_ctx *c = p->ctx;
    CTX_BIND(c);
    ARG_CHECK(offset + n <= p->capacity);
    if (n == 0) return 0;
    photons_join(p);
    char *stage = nullptr;
    CU_CHECK(cudaMallocAsync(&stage, n * 64, c->stream));
    size_t off = 0;
    auto take = [&](const void *want, size_t bytes) -> void * {
        if (!want) return nullptr;
        void *dst = stage + off;
        off += bytes;
        return dst;
    };
    StageOut s{};
    s.pos = (float *)take(pos, n * 12);
    s.dir = (float *)take(dir, n * 12);
    s.pol = (float *)take(pol, n * 12);
    s.wl = (float *)take(wavelengths, n * 4);
    s.t = (float *)take(t, n * 4);
    s.weights = (float *)take(weights, n * 4);
    s.last_hit = (int32_t *)take(last_hit, n * 4);
    s.flags = (uint32_t *)take(flags, n * 4);
    s.evidx = (uint32_t *)take(evidx, n * 4);
    s.channel = (int32_t *)take(channel, n * 4);
    unpack_photons_kernel<<<grid_for(n, 256, c, 8), 256, 0, c->stream>>>(p->v, offset, n, s);
    c->launches++;
    auto get = [&](void *dst, const void *src, size_t bytes) {
        if (dst) cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c->stream);
    };
    get(pos, s.pos, n * 12); get(dir, s.dir, n * 12); get(pol, s.pol, n * 12);
    get(wavelengths, s.wl, n * 4); get(t, s.t, n * 4); get(weights, s.weights, n * 4);
    get(last_hit, s.last_hit, n * 4); get(flags, s.flags, n * 4); get(evidx, s.evidx, n * 4);
    get(channel, s.channel, n * 4);
    cudaError_t e = cudaStreamSynchronize(c->stream);
    cudaFreeAsync(stage, c->stream);
    CU_CHECK(e);
    CU_CHECK(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------
// one field of the packed records <-> a contiguous device array (the GPUArray a reference caller
// hands to its own kernels: chroma/gpu/photon.py:29-46 .pos .dir ... are GPUArrays)
// field: 0 pos 1 dir 2 pol (float[n][3])  3 wavelengths 4 t 7 weights (float[n])
//        5 last_hit_triangles (int32[n])  6 flags 8 evidx 9 channel (uint32[n])
// ---------------------------------------------------------------------------------------
template <bool SET>
__global__ void photon_field_kernel(PhotonView v, uint64_t offset, uint64_t n, int field, uint32_t *a) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const uint64_t j = offset + i;
        if (field <= 2) {
            uint32_t *rec = (uint32_t *)(field == 0 ? v.pt + j : field == 1 ? v.dw + j : v.pw + j);
#pragma unroll
            for (int k = 0; k < 3; k++) {
                if (SET) rec[k] = a[3 * i + k]; else a[3 * i + k] = rec[k];
            }
        } else {
            uint32_t *w = field == 3 ? (uint32_t *)(v.dw + j) + 3 : field == 4 ? (uint32_t *)(v.pt + j) + 3
                        : field == 7 ? (uint32_t *)(v.pw + j) + 3
                        : (uint32_t *)(v.st + j) + (field == 6 ? 0 : field == 5 ? 1 : field == 8 ? 2 : 3);
            if (SET) *w = a[i]; else a[i] = *w;
        }
    }
}

static int photon_field(chroma_photons *p, uint64_t offset, uint64_t n, int field, void *array, bool set) {
    ARG_CHECK(p && (n == 0 || array));
    chroma_ctx *c = p->ctx;
    CTX_BIND(c);
    ARG_CHECK(offset + n <= p->capacity && field >= 0 && field <= 9);
    if (n == 0) return 0;
    TRY_RC(photons_join(p));
    if (set) photon_field_kernel<true><<<grid_for(n, 256, c, 8), 256, 0, c->stream>>>(p->v, offset, n, field, (uint32_t *)array);
    else photon_field_kernel<false><<<grid_for(n, 256, c, 8), 256, 0, c->stream>>>(p->v, offset, n, field, (uint32_t *)array);
    KERNEL_CHECK(c);
    return 0;
}
extern "C" int chroma_photons_field_get(chroma_photons *p, uint64_t offset, uint64_t n, int field, void *out_device) {
    return photon_field(p, offset, n, field, out_device, false);
}
extern "C" int chroma_photons_field_set(chroma_photons *p, uint64_t offset, uint64_t n, int field,
                                        const void *in_device) {
    return photon_field(p, offset, n, field, (void *)in_device, true);
}

extern "C" int chroma_photons_copy(chroma_photons *dst, chroma_photons *src) {
    ARG_CHECK(dst && src && dst->ctx == src->ctx && dst->capacity >= src->capacity);
    chroma_ctx *c = dst->ctx;
    CTX_BIND(c);
    photons_join(dst); photons_join(src);
    size_t nb = src->capacity * 16;
    if (nb == 0) return 0;
    CU_CHECK(cudaMemcpyAsync(dst->v.pt, src->v.pt, nb, cudaMemcpyDeviceToDevice, c->stream));
    CU_CHECK(cudaMemcpyAsync(dst->v.dw, src->v.dw, nb, cudaMemcpyDeviceToDevice, c->stream));
    CU_CHECK(cudaMemcpyAsync(dst->v.pw, src->v.pw, nb, cudaMemcpyDeviceToDevice, c->stream));
    CU_CHECK(cudaMemcpyAsync(dst->v.st, src->v.st, nb, cudaMemcpyDeviceToDevice, c->stream));
    return 0;
}

// ---------------------------------------------------------------------------------------
// photon source on the device (chroma_photons_generate, include/chroma_b200.h)
//
// Replaces host sampling (chroma/sample.py:4-57, chroma/benchmark.py:75-83) + the 60 B/photon
// H2D copy with a write-only stream of the four 16-byte records (64 B/photon, no reads).
// Arithmetic is written with explicit rounding intrinsics (no contraction, IEEE sqrt / divide,
// sincospif instead of the fast-math sin/cos) so the CPU oracle reproduces it to the last bits
// of sincospi.
// ---------------------------------------------------------------------------------------
struct SourceParams {
    float px, py, pz;
    float r[9];
    float cos_half, wl_lo, wl_hi, t0, weight;
    uint32_t evidx, sequence;
    uint64_t seed;
    uint32_t id0;   // global index of photon `offset`
};

__device__ __forceinline__ float3 source_cone(float u0, float u1, float cos_half) {
    float s, c;
    sincospif(__fmul_rn(2.0f, u0), &s, &c);
    const float u = __fmaf_rn(u1, __fsub_rn(1.0f, cos_half), cos_half);
    const float q = __fsqrt_rn(fmaxf(__fsub_rn(1.0f, __fmul_rn(u, u)), 0.0f));
    return f3(__fmul_rn(q, c), __fmul_rn(q, s), u);
}

// MODE 0: streaming stores (evict first), 1: default write-back stores
template <int MODE>
__global__ void __launch_bounds__(256)
generate_photons_kernel(PhotonView v, uint64_t offset, uint64_t n, SourceParams sp) {
    const RngKey K = rng_key(sp.seed, sp.sequence, 3u);
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        Rng g;
        rng_init(g, sp.id0 + (uint32_t)i);
        const float u0 = rng_uniform(g, K), u1 = rng_uniform(g, K), u2 = rng_uniform(g, K),
                    u3 = rng_uniform(g, K), u4 = rng_uniform(g, K);
        const float3 l = source_cone(u0, u1, sp.cos_half);
        float3 d;
        d.x = __fmaf_rn(sp.r[2], l.z, __fmaf_rn(sp.r[1], l.y, __fmul_rn(sp.r[0], l.x)));
        d.y = __fmaf_rn(sp.r[5], l.z, __fmaf_rn(sp.r[4], l.y, __fmul_rn(sp.r[3], l.x)));
        d.z = __fmaf_rn(sp.r[8], l.z, __fmaf_rn(sp.r[7], l.y, __fmul_rn(sp.r[6], l.x)));
        const float3 a = source_cone(u2, u3, -1.0f);
        // cross(a, d), each component as fma(a_j, d_k, -(a_k * d_j))
        float3 e;
        e.x = __fmaf_rn(a.y, d.z, -__fmul_rn(a.z, d.y));
        e.y = __fmaf_rn(a.z, d.x, -__fmul_rn(a.x, d.z));
        e.z = __fmaf_rn(a.x, d.y, -__fmul_rn(a.y, d.x));
        float nn = __fmaf_rn(e.z, e.z, __fmaf_rn(e.y, e.y, __fmul_rn(e.x, e.x)));
        if (!(nn > 1e-12f)) {   // v parallel to dir (measure zero): any perpendicular will do
            e = fabsf(d.z) < 0.9f ? f3(d.y, -d.x, 0.0f) : f3(0.0f, d.z, -d.y);
            nn = __fmaf_rn(e.z, e.z, __fmaf_rn(e.y, e.y, __fmul_rn(e.x, e.x)));
        }
        const float inv = __fdiv_rn(1.0f, __fsqrt_rn(nn));
        const float wl = __fmaf_rn(u4, __fsub_rn(sp.wl_hi, sp.wl_lo), sp.wl_lo);
        const uint64_t j = offset + i;
        const float4 r0 = make_float4(sp.px, sp.py, sp.pz, sp.t0), r1 = make_float4(d.x, d.y, d.z, wl),
                     r2 = make_float4(__fmul_rn(e.x, inv), __fmul_rn(e.y, inv), __fmul_rn(e.z, inv), sp.weight);
        const uint4 r3 = make_uint4(0u, 0xFFFFFFFFu, sp.evidx, 0xFFFFFFFFu);
        if (MODE == 0) {
            __stcs(v.pt + j, r0); __stcs(v.dw + j, r1); __stcs(v.pw + j, r2); __stcs(v.st + j, r3);
        } else {
            v.pt[j] = r0; v.dw[j] = r1; v.pw[j] = r2; v.st[j] = r3;
        }
    }
}

extern "C" int chroma_photons_generate(chroma_photons *p, uint64_t offset, uint64_t n,
                                       const chroma_source *src, uint64_t seed) {
    ARG_CHECK(p && src);
    chroma_ctx *c = p->ctx;
    CTX_BIND(c);
    ARG_CHECK(offset + n <= p->capacity);
    ARG_CHECK(src->cos_half_angle >= -1.0f && src->cos_half_angle <= 1.0f);
    if (n == 0) return 0;
    int r = photons_join(p);   // a staged upload into another window of this container lands first
    if (r) return r;
    SourceParams sp;
    sp.px = src->pos[0]; sp.py = src->pos[1]; sp.pz = src->pos[2];
    for (int k = 0; k < 9; k++) sp.r[k] = src->rot[k];
    sp.cos_half = src->cos_half_angle;
    sp.wl_lo = src->wavelength_lo; sp.wl_hi = src->wavelength_hi;
    sp.t0 = src->t0; sp.weight = src->weight;
    sp.evidx = src->evidx; sp.sequence = src->sequence;
    sp.seed = seed;
    sp.id0 = (uint32_t)(p->id_base + offset);
    const uint32_t per_sm = env_u32("CHROMA_GEN_CTAS_PER_SM", 8);
    if (env_u32("CHROMA_GEN_STORE", 0) == 0)
        generate_photons_kernel<0><<<grid_for(n, 256, c, per_sm), 256, 0, c->stream>>>(p->v, offset, n, sp);
    else
        generate_photons_kernel<1><<<grid_for(n, 256, c, per_sm), 256, 0, c->stream>>>(p->v, offset, n, sp);
    KERNEL_CHECK(c);
    return 0;
}

// propagate.cu:13-52 photon_duplicate
__global__ void duplicate_kernel(PhotonView v, uint64_t n, uint32_t ncopies) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float4 a = v.pt[i], b = v.dw[i], c = v.pw[i];
        uint4 d = v.st[i];
        for (uint32_t k = 1; k < ncopies; k++) {
            uint64_t j = i + (uint64_t)k * n;
            v.pt[j] = a; v.dw[j] = b; v.pw[j] = c; v.st[j] = d;
        }
    }
}

extern "C" int chroma_photons_duplicate(chroma_photons *p, uint64_t n, uint32_t ncopies) {
    ARG_CHECK(p && ncopies >= 1);
    chroma_ctx *c = p->ctx;
    CTX_BIND(c);
    ARG_CHECK(n * ncopies <= p->capacity);
    if (n == 0 || ncopies == 1) return 0;
    photons_join(p);
    duplicate_kernel<<<grid_for(n, 256, c, 8), 256, 0, c->stream>>>(p->v, n, ncopies);
    KERNEL_CHECK(c);
    return 0;
}
