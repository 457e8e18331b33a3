// wide_prop_kernel.cuh — the fused persistent propagate kernel on the wide tree (engine_wide.cu explains it).
// Included by two translation units: engine_wide.cu (kernel `wide_propagate_kernel`, random draws
// inlined: fastest where the search dominates, config C3) and engine_wide_ool.cu (kernel
// `wide_propagate_ool_kernel`, compiled with RNG_DRAW_NOINLINE=1: the draw's buffer logic out of line,
// 10 % less code, fastest where the physics dominates and the instruction cache is the limit,
// config C4: 20.8 -> 17.9 ms per 2.5e6-photon event, r2d).  The launcher picks by what the geometry
// contains (reemitting materials or non-default surface models -> out of line).
#pragma once
#ifndef WPROP_KERNEL_NAME
#define WPROP_KERNEL_NAME wide_propagate_kernel
#endif
#define WPROP_THREADS 256
#ifndef WPROP_MIN_CTAS
#define WPROP_MIN_CTAS 4   // 64 registers: r01v/r01w 3 -> 4 CTAs/SM = +12 %, 5 (48 regs) spills and loses 20 %
#endif

// WPropParams: common.cuh

template <bool STAGED>
__global__ void __launch_bounds__(WPROP_THREADS, WPROP_MIN_CTAS)
WPROP_KERNEL_NAME(GeomView g, WideView w, PhotonView ph, WPropParams prm) {
    extern __shared__ float4 smem_raw[];
    uint2 *s_stack = (uint2 *)smem_raw;
    float *s_opt = (float *)(s_stack + WIDE_SSTACK * WPROP_THREADS);
    for (uint32_t i = threadIdx.x; i < g.opt.smem_floats; i += blockDim.x) s_opt[i] = __ldg(g.opt.blob + i);
    __syncthreads();

    uint2 l_stack[WIDE_STACK - WIDE_SSTACK];
    const WStackRef stk = wstk_make(s_stack + threadIdx.x, l_stack, WPROP_THREADS);
    WSpread spr;
    {
        unsigned char *sb = (unsigned char *)(s_opt + g.opt.smem_floats) + (threadIdx.x >> 5) * WSPREAD_BYTES_PER_WARP;
        spr.res = (float *)sb; spr.own = sb + WSPREAD_PAIRS * 4;
    }
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    const bool use_weights = prm.use_weights != 0;
    const RngKey K = rng_key(prm.seed, prm.epoch, 0u);

    bool have = false, exhausted = false;
    unsigned overflow = 0u;
    uint32_t dummy0 = 0, dummy1 = 0;
    Photon p;
    Rng rng;
    WTrav tv;
    tv.cur = WIDE_NONE; tv.sp = -1; tv.hit = -1; tv.dist = 0.f; tv.last_slot = -1;
    uint64_t idx = 0;
    int steps = 0, sfirst = 0;

    while (true) {
        const bool work = have && !wtrav_done(tv);
        const unsigned mW = __ballot_sync(0xffffffffu, work);
        const unsigned mH = __ballot_sync(0xffffffffu, have);
        if (mH == 0u && exhausted) break;
        const int nidle = __popc(mH & ~mW) + (exhausted ? 0 : 32 - __popc(mH));

        if (mW == 0u || nidle >= prm.refill) {
            // ---------------- physics phase ----------------
            bool start = false;
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
            if (!exhausted) {
                const unsigned hv = __ballot_sync(0xffffffffu, have);
                const int nfree = 32 - __popc(hv);
                if (nfree > 0) {
                    unsigned long long base = 0;
                    if (lane == 0) base = atomicAdd(prm.head, (unsigned long long)nfree);
                    base = __shfl_sync(0xffffffffu, base, 0);
                    if (base + (unsigned long long)nfree >= prm.n) exhausted = true;
                    if (STAGED && base < prm.n) {
                        // wait until the copy engine has delivered the photons this warp just claimed
                        unsigned long long last = base + (unsigned long long)nfree;
                        if (last > prm.n) last = prm.n;
                        const unsigned long long need = prm.offset + last - prm.sv_offset;
                        unsigned stalled = 0u;
                        if (lane == 0) {
                            // watchdog: ~4 s (CHROMA_UPLOAD_WATCHDOG microseconds) without progress raises status bit 1 instead of hanging the GPU
                            unsigned spins = 0;
                            while (*(volatile const unsigned long long *)prm.ready < need) {
                                __nanosleep(1000);
                                if (++spins > prm.watchdog_spins) { atomicOr(prm.status, 2u); stalled = 1u; break; }
                            }
                        }
                        // the upload stalled: what this warp claimed has not landed and is not read (the call
                        // fails with status bit 1; the warp takes no more photons)
                        stalled = __shfl_sync(0xffffffffu, stalled, 0);
                        if (stalled) { exhausted = true; base = prm.n; }
                    }
                    if (!have) {
                        unsigned long long i = base + (unsigned long long)__popc(~hv & lt_mask);
                        if (i < prm.n) {
                            idx = prm.offset + i;
                            uint4 st;
                            float4 a, b, c;
                            if (STAGED) {
                                // host layout -> registers (L2-coherent loads: the data came by DMA)
                                const StageView &sv = prm.sv;
                                const uint64_t k = idx - prm.sv_offset;
                                st = make_uint4(sv.flags ? __ldcg(sv.flags + k) : sv.flags_v,
                                                (uint32_t)(sv.last_hit ? __ldcg(sv.last_hit + k) : sv.last_hit_v),
                                                sv.evidx ? __ldcg(sv.evidx + k) : sv.evidx_v, 0xFFFFFFFFu);
                                a = sv.pos ? make_float4(__ldcg(sv.pos + 3 * k), __ldcg(sv.pos + 3 * k + 1), __ldcg(sv.pos + 3 * k + 2), 0.f)
                                           : make_float4(sv.pos_v[0], sv.pos_v[1], sv.pos_v[2], 0.f);
                                a.w = sv.t ? __ldcg(sv.t + k) : sv.t_v;
                                b = make_float4(__ldcg(sv.dir + 3 * k), __ldcg(sv.dir + 3 * k + 1), __ldcg(sv.dir + 3 * k + 2), __ldcg(sv.wl + k));
                                c = make_float4(__ldcg(sv.pol + 3 * k), __ldcg(sv.pol + 3 * k + 1), __ldcg(sv.pol + 3 * k + 2),
                                                sv.weights ? __ldcg(sv.weights + k) : sv.weight_v);
                                ph.st[idx] = st;   // evidx / channel words of the record
                                if ((st.x & TERMINAL_MASK) != 0) {   // untouched photons still need their records
                                    ph.pt[idx] = a; ph.dw[idx] = b; ph.pw[idx] = c;
                                }
                            } else {
                                st = ph.st[idx];
                                if ((st.x & TERMINAL_MASK) == 0) { a = ph.pt[idx]; b = ph.dw[idx]; c = ph.pw[idx]; }
                            }
                            if ((st.x & TERMINAL_MASK) == 0) {  // propagate.cu:258
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
                if (isnan(p.dir.x * p.dir.y * p.dir.z * p.pos.x * p.pos.y * p.pos.z) ||
                    wide_origin_unsearchable(p.pos, WIDE_GUARD_QUANTA * w.world_scale)) {   // wide.cuh: beyond what the grid can hold
                    p.history |= CHROMA_NO_HIT | (unsigned)CHROMA_NAN_ABORT;
                    atomicAdd(prm.nan_count, 1u);
                    ph.pt[idx] = make_float4(p.pos.x, p.pos.y, p.pos.z, p.time);
                    ph.dw[idx] = make_float4(p.dir.x, p.dir.y, p.dir.z, p.wavelength);
                    ph.pw[idx] = make_float4(p.pol.x, p.pol.y, p.pol.z, p.weight);
                    *reinterpret_cast<uint2 *>(ph.st + idx) = make_uint2(p.history, (uint32_t)p.last_tri);
                    have = false;
                    tv.cur = WIDE_NONE;
                } else {
                    wtrav_begin(tv, w, p.pos, p.dir, p.last_slot);
                }
            }
            continue;
        }
        // ---------------- search phase ----------------
        for (int it = 0; it < prm.iters; it++)
            wtrav_round<false>(tv, have, w, p.pos, p.dir, stk, prm.leaf_thresh, overflow, dummy0, dummy1, spr);
    }
    if (overflow) atomicOr(prm.status, 1u);
}

