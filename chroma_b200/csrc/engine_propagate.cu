// engine_propagate.cu — the persistent propagate kernel, hit compaction, DAQ and PDF.
//
// Replaces chroma/cuda/propagate.cu:217-319 + chroma/gpu/photon.py:193-259 (step loop),
// propagate.cu:54-214 (select / get_flat_hits / copy_queue), chroma/cuda/daq.cu +
// chroma/gpu/daq.py (GPUDaq) and chroma/cuda/pdf.cu:9-32 + chroma/gpu/pdf.py:177-227.
#include <stdlib.h>
#include <cub/device/device_scan.cuh>
#include "physics.cuh"

#define PROP_THREADS 256
#ifndef PROP_MIN_CTAS
#define PROP_MIN_CTAS 2
#endif

#if CHROMA_AB_KERNELS
static uint32_t env_u32(const char *name, uint32_t dflt) {
    const char *s = getenv(name);
    if (!s || !*s) return dflt;
    return (uint32_t)strtoul(s, nullptr, 10);
}
#endif

#if CHROMA_AB_KERNELS
struct PropParams {
    uint64_t offset, n;   // photon range
    uint64_t seed;
    uint32_t id_base;     // RNG id of photon 0 of the container
    uint32_t epoch;
    int max_steps, use_weights, scatter_first;
    int refill;           // refill the warp when >= refill lanes are idle
    int leaf_thresh;      // v2: run the triangle phase when >= leaf_thresh lanes have a parked leaf
    int node_iters;       // v2: node tests per scheduling round
    unsigned long long *head;  // queue head
    uint32_t *nan_count;
    uint32_t *status;     // bit0 = stack overflow
};

// one step of propagate.cu:264-300; returns true when the photon is finished for this call
__device__ __forceinline__ bool propagate_step(Photon &p, Rng &rng, const RngKey &K, const GeomView &g,
                                               const uint4 *s_nodes, const float *s_opt,
                                               unsigned *stack, bool &overflow, bool use_weights,
                                               int &scatter_first, uint32_t *nan_count) {
    // propagate.cu:270-273
    if (isnan(p.dir.x * p.dir.y * p.dir.z * p.pos.x * p.pos.y * p.pos.z)) {
        p.history |= CHROMA_NO_HIT | (unsigned)CHROMA_NAN_ABORT;
        atomicAdd(nan_count, 1u);
        return true;
    }
    State s;
    if (!fill_state(s, p, g, s_nodes, s_opt, stack, overflow)) return true;  // NO_HIT
    int command = propagate_to_boundary(p, s, rng, K, g, s_opt, use_weights, scatter_first);
    scatter_first = 0;  // only once (propagate.cu:281)
    if (command == CMD_BREAK) return true;
    if (command == CMD_CONTINUE) return false;
    if (s.surface_index != -1) {
        command = propagate_at_surface(p, s, rng, K, g, s_opt, use_weights);
        if (command == CMD_BREAK) return true;
        if (command == CMD_CONTINUE) return false;
    }
    propagate_at_boundary(p, s, rng, K);
    return false;
}

__global__ void __launch_bounds__(PROP_THREADS, PROP_MIN_CTAS)
propagate_kernel(GeomView g, PhotonView ph, PropParams prm) {
    extern __shared__ float4 smem_raw[];
    uint4 *s_nodes = (uint4 *)smem_raw;
    float *s_opt = (float *)(s_nodes + g.top_nodes);
    for (uint32_t i = threadIdx.x; i < g.top_nodes; i += blockDim.x) s_nodes[i] = __ldg(g.nodes + i);
    for (uint32_t i = threadIdx.x; i < g.opt.smem_floats; i += blockDim.x) s_opt[i] = __ldg(g.opt.blob + i);
    __syncthreads();

    unsigned stack[CHROMA_STACK_WORDS];
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    const bool use_weights = prm.use_weights != 0;

    bool active = false, exhausted = false, overflow = false;
    Photon p;
    Rng rng;
    const RngKey K = rng_key(prm.seed, prm.epoch, 0u);
    uint64_t idx = 0;
    uint32_t keep_evidx = 0, keep_channel = 0;
    int steps = 0, sfirst = 0;

    while (true) {
        unsigned act = __ballot_sync(0xffffffffu, active);
        int nfree = 32 - __popc(act);
        if (!exhausted && (act == 0u || nfree >= prm.refill)) {
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(prm.head, (unsigned long long)nfree);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (base >= prm.n) {
                exhausted = true;
            } else {
                if (base + (unsigned long long)nfree >= prm.n) exhausted = true;
                if (!active) {
                    unsigned long long i = base + (unsigned long long)__popc(~act & lt_mask);
                    if (i < prm.n) {
                        idx = prm.offset + i;
                        uint4 st = ph.st[idx];
                        // propagate.cu:258: terminal photons are left untouched
                        if ((st.x & TERMINAL_MASK) == 0) {
                            float4 a = ph.pt[idx], b = ph.dw[idx], c = ph.pw[idx];
                            p.pos = f3(a.x, a.y, a.z); p.time = a.w;
                            p.dir = f3(b.x, b.y, b.z); p.wavelength = b.w;
                            p.pol = f3(c.x, c.y, c.z); p.weight = c.w;
                            // propagate.cu:248-250
                            p.dir = p.dir / norm3(p.dir);
                            p.pol = p.pol / norm3(p.pol);
                            p.history = st.x;
                            p.last_tri = (int)st.y;
                            p.last_slot = -1;
                            if (p.last_tri >= 0 && (uint32_t)p.last_tri < g.ntriangles)
                                p.last_slot = g.tri_slot[p.last_tri];
                            keep_evidx = st.z; keep_channel = st.w;
                            rng_init(rng, prm.id_base + (uint32_t)idx);
                            steps = 0;
                            sfirst = prm.scatter_first;
                            active = true;
                        }
                    }
                }
            }
            act = __ballot_sync(0xffffffffu, active);
        }
        if (act == 0u) {
            if (exhausted) break;
            continue;
        }
        if (active) {
            steps++;
            bool done = propagate_step(p, rng, K, g, s_nodes, s_opt, stack, overflow, use_weights,
                                       sfirst, prm.nan_count);
            if (done || steps >= prm.max_steps) {
                ph.pt[idx] = make_float4(p.pos.x, p.pos.y, p.pos.z, p.time);
                ph.dw[idx] = make_float4(p.dir.x, p.dir.y, p.dir.z, p.wavelength);
                ph.pw[idx] = make_float4(p.pol.x, p.pol.y, p.pol.z, p.weight);
                ph.st[idx] = make_uint4(p.history, (uint32_t)p.last_tri, keep_evidx, keep_channel);
                active = false;
            }
        }
    }
    if (overflow) atomicOr(prm.status, 1u);
}


// ---------------------------------------------------------------------------------------
// v2: fused, barrier-free persistent kernel with the warp-scheduled search.
//
// Every lane owns one photon from its first step to its last, so no kernel boundary (and no
// global barrier) separates the steps: the critical path is one photon's history, not the sum
// over steps of the longest search of each step as in the wavefront pair (engine_wavefront.cu).
// The warp votes on what to run:
//   physics phase — when >= `refill` lanes have finished their search (or nothing else can
//                   run): propagate.cu:277-300 for those lanes, retire / start the next
//                   search; empty lanes are refilled from the queue head (one atomicAdd/warp)
//   search phase  — otherwise: `node_iters` rounds of one node test per lane with the
//                   latency-oriented micro-step (traverse.cuh); a leaf whose box passes is
//                   tested at once
// Per photon the sequence of tests, prunes, random draws and physics is unchanged.
// ---------------------------------------------------------------------------------------
#ifndef PROP2_MIN_CTAS
#define PROP2_MIN_CTAS 3
#endif

__global__ void __launch_bounds__(PROP_THREADS, PROP2_MIN_CTAS)
propagate_kernel_v2(GeomView g, PhotonView ph, PropParams prm) {
    extern __shared__ float4 smem_raw[];
    unsigned *s_stack = (unsigned *)smem_raw;
    float *s_opt = (float *)(s_stack + CHROMA_SSTACK * PROP_THREADS);
    for (uint32_t i = threadIdx.x; i < g.opt.smem_floats; i += blockDim.x) s_opt[i] = __ldg(g.opt.blob + i);
    __syncthreads();

    unsigned l_stack[CHROMA_STACK_WORDS - CHROMA_SSTACK];
    const StackRef stk{s_stack + threadIdx.x, l_stack, PROP_THREADS};
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    const bool use_weights = prm.use_weights != 0;
    const RngKey K = rng_key(prm.seed, prm.epoch, 0u);

    bool have = false;       // lane holds a photon
    bool exhausted = false;  // queue drained (warp-uniform)
    unsigned overflow = 0u;
    Photon p;
    Rng rng;
    TravL tv;
    tv.leaf = -1; tv.cur = tv.end = 0; tv.sp = -1;
    uint64_t idx = 0;
    int steps = 0, sfirst = 0;

    while (true) {
        const bool work = have && !travl_done(tv);
        const unsigned mW = __ballot_sync(0xffffffffu, work);
        const unsigned mH = __ballot_sync(0xffffffffu, have);
        if (mH == 0u && exhausted) break;
        const int nidle = __popc(mH & ~mW) + (exhausted ? 0 : 32 - __popc(mH));

        if (mW == 0u || nidle >= prm.refill) {
            // ---------------- physics phase ----------------
            bool start = false;  // lane must start a search
            if (have && !work) {
                bool done;
                State s;
                if (!fill_state_post(s, p, g, s_opt, tv.hit, tv.hit >= 0 ? tv.dist : -1.0f)) {
                    done = true;  // NO_HIT
                } else {
                    int command = propagate_to_boundary(p, s, rng, K, g, s_opt, use_weights, sfirst);
                    sfirst = 0;
                    done = (command == CMD_BREAK);
                    if (command == CMD_PASS) {
                        if (s.surface_index != -1) {
                            command = propagate_at_surface(p, s, rng, K, g, s_opt, use_weights);
                            done = (command == CMD_BREAK);
                        }
                        if (command == CMD_PASS) propagate_at_boundary(p, s, rng, K);
                    }
                }
                if (done || steps >= prm.max_steps) {
                    ph.pt[idx] = make_float4(p.pos.x, p.pos.y, p.pos.z, p.time);
                    ph.dw[idx] = make_float4(p.dir.x, p.dir.y, p.dir.z, p.wavelength);
                    ph.pw[idx] = make_float4(p.pol.x, p.pol.y, p.pol.z, p.weight);
                    *reinterpret_cast<uint2 *>(ph.st + idx) = make_uint2(p.history, (uint32_t)p.last_tri);
                    have = false;
                } else {
                    start = true;
                }
            }
            // refill empty lanes
            if (!exhausted) {
                const unsigned hv = __ballot_sync(0xffffffffu, have);
                const int nfree = 32 - __popc(hv);
                if (nfree > 0) {
                    unsigned long long base = 0;
                    if (lane == 0) base = atomicAdd(prm.head, (unsigned long long)nfree);
                    base = __shfl_sync(0xffffffffu, base, 0);
                    if (base + (unsigned long long)nfree >= prm.n) exhausted = true;
                    if (!have) {
                        unsigned long long i = base + (unsigned long long)__popc(~hv & lt_mask);
                        if (i < prm.n) {
                            idx = prm.offset + i;
                            uint4 st = ph.st[idx];
                            if ((st.x & TERMINAL_MASK) == 0) {  // propagate.cu:258
                                float4 a = ph.pt[idx], b = ph.dw[idx], c = ph.pw[idx];
                                p.pos = f3(a.x, a.y, a.z); p.time = a.w;
                                p.dir = f3(b.x, b.y, b.z); p.wavelength = b.w;
                                p.pol = f3(c.x, c.y, c.z); p.weight = c.w;
                                p.dir = p.dir / norm3(p.dir);  // propagate.cu:248-250
                                p.pol = p.pol / norm3(p.pol);
                                p.history = st.x;
                                p.last_tri = (int)st.y;
                                p.last_slot = -1;
                                if (p.last_tri >= 0 && (uint32_t)p.last_tri < g.ntriangles)
                                    p.last_slot = g.tri_slot[p.last_tri];
                                rng_init(rng, prm.id_base + (uint32_t)idx);
                                steps = 0;
                                sfirst = prm.scatter_first;
                                have = true;
                                start = true;
                            }
                        }
                    }
                }
            }
            if (start) {
                steps++;
                // propagate.cu:270-273
                if (isnan(p.dir.x * p.dir.y * p.dir.z * p.pos.x * p.pos.y * p.pos.z)) {
                    p.history |= CHROMA_NO_HIT | (unsigned)CHROMA_NAN_ABORT;
                    atomicAdd(prm.nan_count, 1u);
                    ph.pt[idx] = make_float4(p.pos.x, p.pos.y, p.pos.z, p.time);
                    ph.dw[idx] = make_float4(p.dir.x, p.dir.y, p.dir.z, p.wavelength);
                    ph.pw[idx] = make_float4(p.pol.x, p.pol.y, p.pol.z, p.weight);
                    *reinterpret_cast<uint2 *>(ph.st + idx) = make_uint2(p.history, (uint32_t)p.last_tri);
                    have = false;
                } else {
                    travl_begin(tv, g, p.pos, p.dir, p.last_slot);
                }
            }
            continue;
        }
        // ---------------- search phase ----------------
        for (int it = 0; it < prm.node_iters; it++)
            if (have && travl_node_work(tv)) travl_step(tv, g, p.pos, p.dir, stk, overflow);
    }
    if (overflow) atomicOr(prm.status, 1u);
}

#endif  // CHROMA_AB_KERNELS

extern "C" int chroma_photons_propagate(chroma_photons *p, chroma_geometry *g, chroma_rng *rng,
                                        uint64_t offset, uint64_t n, int max_steps,
                                        int use_weights, int scatter_first, uint64_t *nan_aborted) {
    ARG_CHECK(p && g && rng);
    chroma_ctx *c = p->ctx;
    CTX_BIND(c);
    ARG_CHECK(g->ctx == c && rng->ctx == c);
    ARG_CHECK(offset + n <= p->capacity);
    if (nan_aborted) *nan_aborted = 0;
    if (n == 0 || max_steps <= 0) return 0;
    // One path in the product library: the fused persistent kernel on the wide tree (engine_wide.cu) or
    // on the two-level tree (engine_inst.cu).  A/B builds (-DCHROMA_AB_KERNELS=1) also carry round 1's
    // kernels on the reference-format tree: CHROMA_PROP_V=3 the wavefront pair (engine_wavefront.cu),
    // 1|2 the fused persistent kernels above.
    const int ver = c->knobs.prop_v;
    if (g->instanced || ver >= 4 || !CHROMA_AB_KERNELS)
        return chroma_wide_propagate(p, g, rng, offset, n, max_steps, use_weights, scatter_first, nan_aborted);
#if CHROMA_AB_KERNELS
    photons_join(p);
    if (ver >= 3)
        return chroma_wave_propagate(p, g, rng, offset, n, max_steps, use_weights, scatter_first, nan_aborted);
    const bool v1 = ver == 1;
    size_t smem = (v1 ? (size_t)g->view.top_nodes * sizeof(uint4) : (size_t)CHROMA_SSTACK * PROP_THREADS * sizeof(unsigned)) +
                  (size_t)g->view.opt.smem_floats * sizeof(float);
    static bool attr_set = false;
    if (!attr_set) {
        CU_CHECK(cudaFuncSetAttribute(propagate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_set = true;
    }
    // scratch words: [4..5] queue head (u64), [6] nan count, [7] status
    CU_CHECK(cudaMemsetAsync(c->d_scratch + 4, 0, 16, c->stream));
    PropParams prm;
    prm.offset = offset; prm.n = n;
    prm.seed = rng->seed;
    prm.id_base = (uint32_t)p->id_base;
    prm.epoch = rng->epoch++;
    prm.max_steps = max_steps; prm.use_weights = use_weights; prm.scatter_first = scatter_first;
    prm.refill = (int)env_u32("CHROMA_REFILL", 8);
    if (prm.refill < 1) prm.refill = 1;
    if (prm.refill > 32) prm.refill = 32;
    prm.head = (unsigned long long *)(c->d_scratch + 4);
    prm.nan_count = c->d_scratch + 6;
    prm.status = c->d_scratch + 7;
    prm.leaf_thresh = (int)env_u32("CHROMA_LEAF_THRESH", 4);
    prm.node_iters = (int)env_u32("CHROMA_NODE_ITERS", 8);
    if (prm.leaf_thresh < 1) prm.leaf_thresh = 1;
    if (prm.node_iters < 1) prm.node_iters = 1;
    static bool attr2_set = false;
    if (!attr2_set) {
        CU_CHECK(cudaFuncSetAttribute(propagate_kernel_v2, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr2_set = true;
    }
    int occ = 0;
    if (v1) CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, propagate_kernel, PROP_THREADS, smem));
    else CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, propagate_kernel_v2, PROP_THREADS, smem));
    if (occ < 1) occ = 1;
    unsigned grid = grid_for(n, PROP_THREADS, c, (unsigned)occ);
    if (v1) propagate_kernel<<<grid, PROP_THREADS, smem, c->stream>>>(g->view, p->v, prm);
    else propagate_kernel_v2<<<grid, PROP_THREADS, smem, c->stream>>>(g->view, p->v, prm);
    KERNEL_CHECK(c);
    if (nan_aborted) {
        CU_CHECK(cudaMemcpyAsync(c->h_scratch + 4, c->d_scratch + 4, 16, cudaMemcpyDeviceToHost, c->stream));
        CU_CHECK(cudaStreamSynchronize(c->stream));
        *nan_aborted = c->h_scratch[6];
        if (c->h_scratch[7] & 1u) {
            chroma_set_error("BVH traversal stack overflow (> %d pending sibling groups)", CHROMA_STACK_WORDS);
            return CHROMA_ERR_OVERFLOW;
        }
    }
    return 0;
#else
    return CHROMA_ERR_STATE;   // not reached: the branch above always returns in the product library
#endif  // CHROMA_AB_KERNELS
}

// ---------------------------------------------------------------------------------------
// stream compaction (select / flat hits / live ids): deterministic, ascending photon order
//   pass 1: per-block match count (warp ballot + popc)   pass 2: exclusive scan of counts
//   pass 3: scatter with ballot ranks
// ---------------------------------------------------------------------------------------
enum { MATCH_FLAG = 0, MATCH_HIT = 1, MATCH_LIVE = 2 };

struct MatchArgs {
    int mode;
    uint32_t flag;
    const int32_t *tri_solid_slot;  // unused
    const float4 *tris;             // packed triangle records (solid id in c.w)
    const int32_t *tri_slot;
    uint32_t ntriangles;
    const uint32_t *inst_tri_base;  // two-level geometry: first triangle of every solid, else NULL
    uint32_t ninst;
    const int32_t *solid_to_channel;
    uint64_t nsolids;
};

__device__ __forceinline__ bool photon_matches(const MatchArgs &m, const uint4 &st, int &channel) {
    channel = -1;
    if (m.mode == MATCH_LIVE) return (st.x & TERMINAL_MASK) == 0;
    if (!(st.x & m.flag)) return false;
    if (m.mode == MATCH_FLAG) return true;
    // propagate.cu:160-167: detected on a triangle whose solid maps to a channel
    uint32_t solid;
    if (!solid_of_triangle(m.tris, m.tri_slot, m.inst_tri_base, m.ninst, m.ntriangles, (int)st.y, solid)) return false;
    if (solid >= m.nsolids) return false;
    channel = m.solid_to_channel[solid];
    return channel >= 0;
}

#define COMPACT_THREADS 256

__global__ void __launch_bounds__(COMPACT_THREADS)
match_count_kernel(PhotonView v, uint64_t offset, uint64_t n, MatchArgs m, uint32_t *block_counts) {
    __shared__ uint32_t warp_counts[COMPACT_THREADS / 32];
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool match = false;
    int ch;
    if (i < n) match = photon_matches(m, v.st[offset + i], ch);
    unsigned b = __ballot_sync(0xffffffffu, match);
    if ((threadIdx.x & 31) == 0) warp_counts[threadIdx.x >> 5] = __popc(b);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t s = 0;
        for (int w = 0; w < COMPACT_THREADS / 32; w++) s += warp_counts[w];
        block_counts[blockIdx.x] = s;
    }
}

__global__ void __launch_bounds__(COMPACT_THREADS)
match_scatter_kernel(PhotonView v, uint64_t offset, uint64_t n, MatchArgs m,
                     const uint32_t *block_offsets, PhotonView out, uint32_t *out_ids) {
    __shared__ uint32_t warp_counts[COMPACT_THREADS / 32];
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool match = false;
    int ch = -1;
    uint4 st = make_uint4(0, 0, 0, 0);
    if (i < n) {
        st = v.st[offset + i];
        match = photon_matches(m, st, ch);
    }
    unsigned b = __ballot_sync(0xffffffffu, match);
    unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) warp_counts[warp] = __popc(b);
    __syncthreads();
    if (match) {
        uint32_t pos = block_offsets[blockIdx.x];
        for (unsigned w = 0; w < warp; w++) pos += warp_counts[w];
        pos += __popc(b & ((1u << lane) - 1u));
        if (out_ids) {
            out_ids[pos] = (uint32_t)(offset + i);
        } else {
            out.pt[pos] = v.pt[offset + i];
            out.dw[pos] = v.dw[offset + i];
            out.pw[pos] = v.pw[offset + i];
            st.w = (uint32_t)ch;
            out.st[pos] = st;
        }
    }
}

// runs the three passes; returns the match count and (optionally) a new container / id list
static int compact(chroma_photons *p, uint64_t offset, uint64_t n, const MatchArgs &m,
                   chroma_photons **out, uint32_t *ids_host, uint64_t *ids_count) {
    chroma_ctx *c = p->ctx;
    photons_join(p);
    uint64_t nblocks = (n + COMPACT_THREADS - 1) / COMPACT_THREADS;
    uint64_t total = 0;
    uint32_t *d_counts = nullptr, *d_offsets = nullptr, *d_ids = nullptr;
    void *d_temp = nullptr;
    int rc = 0;
    chroma_photons *res = nullptr;
    do {
        if (n > 0) {
            if (cudaMallocAsync(&d_counts, nblocks * 4, c->stream) || cudaMallocAsync(&d_offsets, nblocks * 4, c->stream)) {
                chroma_set_error("cudaMalloc: %s", cudaGetErrorString(cudaGetLastError()));
                rc = CHROMA_ERR_CUDA; break;
            }
            match_count_kernel<<<(unsigned)nblocks, COMPACT_THREADS, 0, c->stream>>>(p->v, offset, n, m, d_counts);
            c->launches++;
            size_t tb = 0;
            cub::DeviceScan::ExclusiveSum(nullptr, tb, d_counts, d_offsets, (int)nblocks, c->stream);
            if (cudaMallocAsync(&d_temp, tb ? tb : 16, c->stream)) { chroma_set_error("cudaMalloc scan temp"); rc = CHROMA_ERR_CUDA; break; }
            cub::DeviceScan::ExclusiveSum(d_temp, tb, d_counts, d_offsets, (int)nblocks, c->stream);
            c->launches++;
            uint32_t lc = 0, lo = 0;
            cudaMemcpyAsync(&lc, d_counts + nblocks - 1, 4, cudaMemcpyDeviceToHost, c->stream);
            cudaMemcpyAsync(&lo, d_offsets + nblocks - 1, 4, cudaMemcpyDeviceToHost, c->stream);
            if (cudaStreamSynchronize(c->stream) != cudaSuccess) {
                chroma_set_error("compaction count: %s", cudaGetErrorString(cudaGetLastError()));
                rc = CHROMA_ERR_CUDA; break;
            }
            total = (uint64_t)lc + lo;
        }
        if (out) {
            if ((rc = chroma_photons_create(c, total, p->id_base, &res))) break;
            res->has_channel = (m.mode == MATCH_HIT);
        }
        if (ids_host) {
            if (total > *ids_count) { chroma_set_error("id buffer too small (%llu > %llu)", (unsigned long long)total, (unsigned long long)*ids_count); rc = CHROMA_ERR_ARG; break; }
            if (total && cudaMallocAsync(&d_ids, total * 4, c->stream)) { chroma_set_error("cudaMalloc ids"); rc = CHROMA_ERR_CUDA; break; }
        }
        if (total > 0) {
            PhotonView ov = res ? res->v : PhotonView{};
            match_scatter_kernel<<<(unsigned)nblocks, COMPACT_THREADS, 0, c->stream>>>(p->v, offset, n, m, d_offsets, ov, d_ids);
            c->launches++;
            if (ids_host) cudaMemcpyAsync(ids_host, d_ids, total * 4, cudaMemcpyDeviceToHost, c->stream);
            if (cudaStreamSynchronize(c->stream) != cudaSuccess) {
                chroma_set_error("compaction scatter: %s", cudaGetErrorString(cudaGetLastError()));
                rc = CHROMA_ERR_CUDA; break;
            }
        }
        if (ids_count) *ids_count = total;
    } while (0);
    if (d_counts) cudaFreeAsync(d_counts, c->stream);
    if (d_offsets) cudaFreeAsync(d_offsets, c->stream);
    if (d_temp) cudaFreeAsync(d_temp, c->stream);
    if (d_ids) cudaFreeAsync(d_ids, c->stream);
    if (rc) {
        if (res) chroma_photons_destroy(res);
        return rc;
    }
    if (out) *out = res;
    return 0;
}

extern "C" int chroma_photons_select(chroma_photons *p, uint32_t target_flag, uint64_t offset,
                                     uint64_t n, chroma_photons **out) {
    ARG_CHECK(p && out);
    CTX_BIND(p->ctx);
    ARG_CHECK(offset + n <= p->capacity);
    MatchArgs m{};
    m.mode = MATCH_FLAG;
    m.flag = target_flag;
    return compact(p, offset, n, m, out, nullptr, nullptr);
}

extern "C" int chroma_photons_flat_hits(chroma_photons *p, chroma_geometry *g, uint32_t target_flag,
                                        uint64_t offset, uint64_t n, chroma_photons **out) {
    ARG_CHECK(p && g && out);
    CTX_BIND(p->ctx);
    ARG_CHECK(offset + n <= p->capacity);
    if (!g->has_detector) {
        chroma_set_error("geometry has no detector attached (GPUDetector required)");
        return CHROMA_ERR_STATE;
    }
    MatchArgs m{};
    m.mode = MATCH_HIT;
    m.flag = target_flag;
    m.tris = g->d_tris;
    m.tri_slot = g->d_tri_slot;
    m.ntriangles = (uint32_t)g->ntriangles;
    m.inst_tri_base = g->view.inst_tri_base;
    m.ninst = g->view.ninst;
    m.solid_to_channel = g->d_solid_to_channel;
    m.nsolids = g->nsolids;
    return compact(p, offset, n, m, out, nullptr, nullptr);
}

extern "C" int chroma_photons_live_ids(chroma_photons *p, uint64_t offset, uint64_t n,
                                       uint32_t *ids_host, uint64_t *count) {
    ARG_CHECK(p && ids_host && count);
    CTX_BIND(p->ctx);
    ARG_CHECK(offset + n <= p->capacity);
    MatchArgs m{};
    m.mode = MATCH_LIVE;
    return compact(p, offset, n, m, nullptr, ids_host, count);
}

// propagate.cu:116-144 copy_photon_queue
__global__ void gather_kernel(PhotonView v, const uint32_t *ids, uint64_t n, uint64_t cap, PhotonView out) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        uint64_t j = ids[i];
        if (j >= cap) j = cap - 1;
        out.pt[i] = v.pt[j]; out.dw[i] = v.dw[j]; out.pw[i] = v.pw[j]; out.st[i] = v.st[j];
    }
}

extern "C" int chroma_photons_gather(chroma_photons *p, const uint32_t *ids_host, uint64_t n,
                                     chroma_photons **out) {
    ARG_CHECK(p && out && (n == 0 || ids_host));
    chroma_ctx *c = p->ctx;
    CTX_BIND(c);
    chroma_photons *res = nullptr;
    photons_join(p);
    int rc = chroma_photons_create(c, n, p->id_base, &res);
    if (rc) return rc;
    if (n > 0) {
        ARG_CHECK(p->capacity > 0);
        uint32_t *d_ids = nullptr;
        if (cudaMallocAsync(&d_ids, n * 4, c->stream)) { chroma_photons_destroy(res); chroma_set_error("cudaMalloc ids"); return CHROMA_ERR_CUDA; }
        cudaMemcpyAsync(d_ids, ids_host, n * 4, cudaMemcpyHostToDevice, c->stream);
        gather_kernel<<<grid_for(n, 256, c, 8), 256, 0, c->stream>>>(p->v, d_ids, n, p->capacity, res->v);
        c->launches++;
        cudaError_t e = cudaStreamSynchronize(c->stream);
        cudaFreeAsync(d_ids, c->stream);
        if (e != cudaSuccess) { chroma_photons_destroy(res); chroma_set_error("gather: %s", cudaGetErrorString(e)); return CHROMA_ERR_CUDA; }
    }
    *out = res;
    return 0;
}
