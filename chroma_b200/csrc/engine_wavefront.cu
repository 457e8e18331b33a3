// engine_wavefront.cu — wavefront form of the propagate step loop (default engine path).
//
// Replaces chroma/gpu/photon.py:193-259 (host step loop with queue swap) and
// chroma/cuda/propagate.cu:217-319 (per-thread step loop).  One propagation step is two
// stream-ordered kernels over the queue of live photons; nothing returns to the host between
// steps (the reference reads the queue length back after every step):
//
//   trace_kernel    persistent, warp-scheduled BVH search (traverse.cuh).  A lane pulls the
//                   next live photon from the queue, reads only its ray (pos, dir, excluded
//                   leaf) and writes (hit leaf, distance).  No physics state is live, so the
//                   kernel needs ~half the registers of a fused kernel and twice the warps
//                   hide the dependent node fetches; lanes that finish are refilled at once.
//   physics_kernel  one thread per live photon, all lanes busy: second half of fill_state,
//                   propagate_to_boundary, surface / boundary physics (physics.cuh), state
//                   write-back, and warp-ballot compaction of the survivors into the next
//                   queue (one atomicAdd per warp).
//
// Per photon the sequence of box tests, triangle tests, random draws and physics equals the
// reference's; only the interleaving between photons differs.  The random stream is
// counter-based (photon id, call epoch), so a photon's draws do not depend on which lane or
// step kernel processes it; only the count of words drawn is carried between steps.
#include "common.cuh"
#if CHROMA_AB_KERNELS
#include <stdlib.h>
#include "physics.cuh"

#define TRACE_THREADS 256
#define PHYS_THREADS 128
#ifndef TRACE_MIN_CTAS
#define TRACE_MIN_CTAS 4
#endif

static uint32_t env_u32w(const char *name, uint32_t dflt) {
    const char *s = getenv(name);
    if (!s || !*s) return dflt;
    return (uint32_t)strtoul(s, nullptr, 10);
}

struct WaveParams {
    uint64_t offset, n;          // photon range of the call
    const uint32_t *queue;       // live ids of this step (NULL: identity over [offset, offset+n))
    uint32_t *next_queue;        // survivors (physics kernel)
    unsigned long long *ctr;     // [0..3] queue lengths, [4..7] fetch heads, by step mod 4
    uint2 *aux;
    int2 *hit;
    int step, max_steps;
    uint64_t seed;
    uint32_t id_base, epoch;
    int use_weights, scatter_first;
    int refill, node_iters, leaf_thresh;
    uint32_t *nan_count, *status;
};

__device__ __forceinline__ unsigned long long wave_count(const WaveParams &w) {
    return w.queue ? w.ctr[w.step & 3] : w.n;
}

// ---------------------------------------------------------------------------------------
// trace: (pos, dir, excluded leaf) -> (hit leaf, distance)
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TRACE_THREADS, TRACE_MIN_CTAS)
trace_kernel(GeomView g, PhotonView ph, WaveParams w) {
    extern __shared__ unsigned s_stack[];
    unsigned l_stack[CHROMA_STACK_WORDS - CHROMA_SSTACK];
    const StackRef stk{s_stack + threadIdx.x, l_stack, TRACE_THREADS};
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    const unsigned long long count = wave_count(w);
    unsigned long long *head = w.ctr + 4 + (w.step & 3);
    bool have = false, exhausted = false;
    unsigned overflow = 0u;
    float3 o = f3(0, 0, 0), d = f3(0, 0, 1);
    uint32_t idx = 0;
    TravL tv;
    tv.leaf = -1; tv.cur = tv.end = 0; tv.sp = -1;

    while (true) {
        const bool work = have && !travl_done(tv);
        const unsigned mW = __ballot_sync(0xffffffffu, work);
        const unsigned mH = __ballot_sync(0xffffffffu, have);
        if (mH == 0u && exhausted) break;
        const int nidle = __popc(mH & ~mW) + (exhausted ? 0 : 32 - __popc(mH));
        if (mW == 0u || nidle >= w.refill) {
            // ---- emit finished searches, pull new rays ----
            if (have && !work) {
                w.hit[idx] = make_int2(tv.hit, tv.hit >= 0 ? __float_as_int(tv.dist) : __float_as_int(-1.0f));
                have = false;
            }
            if (!exhausted) {
                const unsigned hv = __ballot_sync(0xffffffffu, have);
                const int nfree = 32 - __popc(hv);
                unsigned long long base = 0;
                if (lane == 0) base = atomicAdd(head, (unsigned long long)nfree);
                base = __shfl_sync(0xffffffffu, base, 0);
                if (base + (unsigned long long)nfree >= count) exhausted = true;
                if (!have) {
                    unsigned long long e = base + (unsigned long long)__popc(~hv & lt_mask);
                    if (e < count) {
                        idx = w.queue ? w.queue[e] : (uint32_t)(w.offset + e);
                        bool live = true;
                        if (!w.queue) live = (ph.st[idx].x & TERMINAL_MASK) == 0;  // propagate.cu:258
                        if (live) {
                            float4 a = ph.pt[idx], b = ph.dw[idx];
                            o = f3(a.x, a.y, a.z);
                            d = f3(b.x, b.y, b.z);
                            // propagate.cu:270-273: a NaN ray is not traced (it would pass every box)
                            if (isnan(d.x * d.y * d.z * o.x * o.y * o.z)) {
                                w.hit[idx] = make_int2(-2, 0);
                            } else {
                                travl_begin(tv, g, o, d, (int)w.aux[idx].x);
                                have = true;
                            }
                        }
                    }
                }
            }
            continue;
        }
        // ---- search: `node_iters` rounds of one micro-step per lane (box test; a leaf whose box
        // passes is tested in the same round)
        for (int it = 0; it < w.node_iters; it++)
            if (have && travl_node_work(tv)) travl_step(tv, g, o, d, stk, overflow);
    }
    if (overflow) atomicOr(w.status, 1u);
}

// ---------------------------------------------------------------------------------------
// first call step: renormalise (propagate.cu:248-250), resolve the excluded leaf
// ---------------------------------------------------------------------------------------
__global__ void wave_init_kernel(GeomView g, PhotonView ph, WaveParams w) {
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < w.n; e += stride) {
        const uint64_t idx = w.offset + e;
        uint4 st = ph.st[idx];
        if (st.x & TERMINAL_MASK) continue;
        float4 b = ph.dw[idx], c = ph.pw[idx];
        float3 dir = f3(b.x, b.y, b.z), pol = f3(c.x, c.y, c.z);
        dir = dir / norm3(dir);
        pol = pol / norm3(pol);
        ph.dw[idx] = make_float4(dir.x, dir.y, dir.z, b.w);
        ph.pw[idx] = make_float4(pol.x, pol.y, pol.z, c.w);
        int last_tri = (int)st.y, last_slot = -1;
        if (last_tri >= 0 && (uint32_t)last_tri < g.ntriangles) last_slot = g.tri_slot[last_tri];
        w.aux[idx] = make_uint2((uint32_t)last_slot, 0u);
    }
}

// ---------------------------------------------------------------------------------------
// physics: propagate.cu:277-300 for every live photon of the step
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(PHYS_THREADS)
physics_kernel(GeomView g, PhotonView ph, WaveParams w) {
    extern __shared__ float4 smem_raw[];
    float *s_opt = (float *)smem_raw;
    const unsigned long long count = wave_count(w);
    // thread 0 recycles the counters two steps ahead (free: step-2 kernels have completed)
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        w.ctr[(w.step + 2) & 3] = 0ull;
        w.ctr[4 + ((w.step + 2) & 3)] = 0ull;
    }
    if ((unsigned long long)blockIdx.x * blockDim.x >= count) return;  // nothing for this block
    for (uint32_t i = threadIdx.x; i < g.opt.smem_floats; i += blockDim.x) s_opt[i] = __ldg(g.opt.blob + i);
    __syncthreads();
    const unsigned lane = threadIdx.x & 31u;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const RngKey K = rng_key(w.seed, w.epoch, 0u);
    const bool use_weights = w.use_weights != 0;
    const uint64_t rounds = (count + stride - 1) / stride;
    for (uint64_t r = 0; r < rounds; r++) {
        const uint64_t e = r * stride + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
        bool alive = false;
        uint32_t idx = 0;
        if (e < count) {
            idx = w.queue ? w.queue[e] : (uint32_t)(w.offset + e);
            uint2 stxy = *reinterpret_cast<const uint2 *>(ph.st + idx);
            if (w.queue || (stxy.x & TERMINAL_MASK) == 0) {
                Photon p;
                float4 a = ph.pt[idx], b = ph.dw[idx], c = ph.pw[idx];
                p.pos = f3(a.x, a.y, a.z); p.time = a.w;
                p.dir = f3(b.x, b.y, b.z); p.wavelength = b.w;
                p.pol = f3(c.x, c.y, c.z); p.weight = c.w;
                p.history = stxy.x;
                p.last_tri = (int)stxy.y;
                const uint2 ax = w.aux[idx];
                p.last_slot = (int)ax.x;
                const int2 h = w.hit[idx];
                Rng rng;
                rng_seek(rng, K, w.id_base + idx, ax.y);
                bool done;
                if (h.x == -2) {
                    p.history |= CHROMA_NO_HIT | (unsigned)CHROMA_NAN_ABORT;
                    atomicAdd(w.nan_count, 1u);
                    done = true;
                } else {
                    State s;
                    if (!fill_state_post(s, p, g, s_opt, h.x, __int_as_float(h.y))) {
                        done = true;  // NO_HIT
                    } else {
                        int sfirst = (w.step == 1) ? w.scatter_first : 0;  // propagate.cu:281
                        int command = propagate_to_boundary(p, s, rng, K, g, s_opt, use_weights, sfirst);
                        done = (command == CMD_BREAK);
                        if (command == CMD_PASS) {
                            if (s.surface_index != -1) {
                                command = propagate_at_surface(p, s, rng, K, g, s_opt, use_weights);
                                done = (command == CMD_BREAK);
                            }
                            if (command == CMD_PASS) propagate_at_boundary(p, s, rng, K);
                        }
                    }
                }
                ph.pt[idx] = make_float4(p.pos.x, p.pos.y, p.pos.z, p.time);
                ph.dw[idx] = make_float4(p.dir.x, p.dir.y, p.dir.z, p.wavelength);
                ph.pw[idx] = make_float4(p.pol.x, p.pol.y, p.pol.z, p.weight);
                *reinterpret_cast<uint2 *>(ph.st + idx) = make_uint2(p.history, (uint32_t)p.last_tri);
                alive = !done && w.step < w.max_steps;
                if (alive) w.aux[idx] = make_uint2((uint32_t)p.last_slot, rng_tell(rng));
            }
        }
        // warp-ballot compaction of the survivors into the next queue
        const unsigned m = __ballot_sync(0xffffffffu, alive);
        if (m) {
            unsigned long long base = 0;
            if (lane == (unsigned)(__ffs(m) - 1)) base = atomicAdd(w.ctr + ((w.step + 1) & 3), (unsigned long long)__popc(m));
            base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
            if (alive) w.next_queue[base + __popc(m & ((1u << lane) - 1u))] = idx;
        }
    }
}

// ---------------------------------------------------------------------------------------
// host driver
// ---------------------------------------------------------------------------------------
#define WAVE_MAX_CHUNKS 16

static int wave_reserve(chroma_photons *p) {
    WaveScratch &ws = p->wave;
    if (ws.capacity >= p->capacity && ws.aux) return 0;
    if (ws.aux) cudaFree(ws.aux);
    if (ws.hit) cudaFree(ws.hit);
    if (ws.queue[0]) cudaFree(ws.queue[0]);
    if (ws.queue[1]) cudaFree(ws.queue[1]);
    if (ws.ctr) cudaFree(ws.ctr);
    ws = WaveScratch();
    const size_t cap = p->capacity ? p->capacity : 1;
    CU_CHECK(cudaMalloc(&ws.aux, cap * sizeof(uint2)));
    CU_CHECK(cudaMalloc(&ws.hit, cap * sizeof(int2)));
    CU_CHECK(cudaMalloc(&ws.queue[0], cap * sizeof(uint32_t)));
    CU_CHECK(cudaMalloc(&ws.queue[1], cap * sizeof(uint32_t)));
    CU_CHECK(cudaMalloc(&ws.ctr, 8 * WAVE_MAX_CHUNKS * sizeof(unsigned long long)));
    ws.capacity = p->capacity;
    return 0;
}

static int wave_streams(chroma_ctx *c, int k) {
    while (c->n_wave_streams < k) {
        CU_CHECK(cudaStreamCreateWithFlags(&c->wave_stream[c->n_wave_streams], cudaStreamNonBlocking));
        CU_CHECK(cudaEventCreateWithFlags(&c->wave_event[c->n_wave_streams], cudaEventDisableTiming));
        c->n_wave_streams++;
    }
    if (!c->wave_fork) CU_CHECK(cudaEventCreateWithFlags(&c->wave_fork, cudaEventDisableTiming));
    return 0;
}

// The photon range is cut into independent chunks, each with its own queues and its own
// stream: a step kernel only fills the machine while its queue lasts, and the tail of one
// chunk's step (a few long searches) overlaps the next chunk's kernels instead of idling SMs.
int chroma_wave_propagate(chroma_photons *p, chroma_geometry *g, chroma_rng *rng, uint64_t offset,
                          uint64_t n, int max_steps, int use_weights, int scatter_first,
                          uint64_t *nan_aborted) {
    chroma_ctx *c = p->ctx;
    if (p->capacity >= (1ull << 32)) {
        chroma_set_error("wavefront propagate: container too large for 32-bit photon ids");
        return CHROMA_ERR_ARG;
    }
    int rc = wave_reserve(p);
    if (rc) return rc;
    WaveScratch &ws = p->wave;
    static bool attr_set = false;
    if (!attr_set) {
        CU_CHECK(cudaFuncSetAttribute(physics_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        attr_set = true;
    }
    int nchunks = (int)env_u32w("CHROMA_WAVE_CHUNKS", 1);
    const uint64_t min_chunk = env_u32w("CHROMA_WAVE_MIN_CHUNK", 131072);
    if (nchunks > WAVE_MAX_CHUNKS) nchunks = WAVE_MAX_CHUNKS;
    while (nchunks > 1 && n / nchunks < min_chunk) nchunks--;
    if (nchunks < 1) nchunks = 1;
    if ((rc = wave_streams(c, nchunks))) return rc;

    CU_CHECK(cudaMemsetAsync(ws.ctr, 0, 8 * WAVE_MAX_CHUNKS * sizeof(unsigned long long), c->stream));
    CU_CHECK(cudaMemsetAsync(c->d_scratch + 6, 0, 8, c->stream));  // [6] nan count, [7] status

    WaveParams base;
    base.aux = ws.aux; base.hit = ws.hit;
    base.max_steps = max_steps;
    base.seed = rng->seed;
    base.id_base = (uint32_t)p->id_base;
    base.epoch = rng->epoch++;
    base.use_weights = use_weights; base.scatter_first = scatter_first;
    base.refill = (int)env_u32w("CHROMA_TRACE_REFILL", 4);
    base.node_iters = (int)env_u32w("CHROMA_TRACE_ITERS", 8);
    base.leaf_thresh = (int)env_u32w("CHROMA_LEAF_THRESH", 1);
    if (base.refill < 1) base.refill = 1;
    if (base.node_iters < 1) base.node_iters = 1;
    if (base.leaf_thresh < 1) base.leaf_thresh = 1;
    base.nan_count = c->d_scratch + 6;
    base.status = c->d_scratch + 7;

    int occ_t = 0, occ_p = 0;
    const size_t smem_p = (size_t)g->view.opt.smem_floats * sizeof(float);
    const size_t smem_t = (size_t)CHROMA_SSTACK * TRACE_THREADS * sizeof(unsigned);
    CU_CHECK(cudaFuncSetAttribute(trace_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_t));
    CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_t, trace_kernel, TRACE_THREADS, smem_t));
    CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_p, physics_kernel, PHYS_THREADS, smem_p));
    if (occ_t < 1) occ_t = 1;
    if (occ_p < 1) occ_p = 1;

    // fork: the chunk streams start after everything already queued on the context stream
    CU_CHECK(cudaEventRecord(c->wave_fork, c->stream));
    uint64_t c_off[WAVE_MAX_CHUNKS], c_n[WAVE_MAX_CHUNKS], live[WAVE_MAX_CHUNKS];
    for (int k = 0; k < nchunks; k++) {
        c_off[k] = offset + n * (uint64_t)k / nchunks;
        c_n[k] = offset + n * (uint64_t)(k + 1) / nchunks - c_off[k];
        live[k] = c_n[k];
        CU_CHECK(cudaStreamWaitEvent(c->wave_stream[k], c->wave_fork, 0));
    }
    const int check_every = (int)env_u32w("CHROMA_WAVE_CHECK", 4);
    unsigned long long *h_live = (unsigned long long *)(c->h_scratch + 16);  // 16 x u64, pinned
    int step = 1;
    bool any = true;
    while (step <= max_steps && any) {
        const int batch_end = (check_every > 0 && step + check_every - 1 < max_steps) ? step + check_every - 1 : max_steps;
        for (int k = 0; k < nchunks; k++) {
            if (live[k] == 0) continue;
            cudaStream_t st = c->wave_stream[k];
            WaveParams w = base;
            w.offset = c_off[k]; w.n = c_n[k];
            w.ctr = ws.ctr + 8 * k;
            // a chunk's survivors never outnumber the chunk: its queue slice starts at its offset
            uint32_t *q0 = ws.queue[0] + c_off[k], *q1 = ws.queue[1] + c_off[k];
            for (int s = step; s <= batch_end; s++) {
                w.step = s;
                if (s == 1) {
                    wave_init_kernel<<<grid_for(c_n[k], 256, c, 8), 256, 0, st>>>(g->view, p->v, w);
                    c->launches++;
                }
                w.queue = (s == 1) ? nullptr : ((s & 1) ? q1 : q0);
                w.next_queue = ((s + 1) & 1) ? q1 : q0;
                trace_kernel<<<grid_for(live[k], TRACE_THREADS, c, (unsigned)occ_t), TRACE_THREADS, smem_t, st>>>(g->view, p->v, w);
                physics_kernel<<<grid_for(live[k], PHYS_THREADS, c, (unsigned)occ_p), PHYS_THREADS, smem_p, st>>>(g->view, p->v, w);
                c->launches += 2;
            }
            if (batch_end < max_steps)
                CU_CHECK(cudaMemcpyAsync(h_live + k, ws.ctr + 8 * k + ((batch_end + 1) & 3), 8, cudaMemcpyDeviceToHost, st));
        }
        CU_CHECK(cudaGetLastError());
        step = batch_end + 1;
        if (step <= max_steps) {
            // the only host round trip: bound the remaining launches by the survivor counts
            any = false;
            for (int k = 0; k < nchunks; k++) {
                if (live[k] == 0) continue;
                CU_CHECK(cudaStreamSynchronize(c->wave_stream[k]));
                live[k] = h_live[k];
                if (live[k]) any = true;
            }
        }
    }
    // join
    for (int k = 0; k < nchunks; k++) {
        CU_CHECK(cudaEventRecord(c->wave_event[k], c->wave_stream[k]));
        CU_CHECK(cudaStreamWaitEvent(c->stream, c->wave_event[k], 0));
    }
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
}

#endif  // CHROMA_AB_KERNELS
