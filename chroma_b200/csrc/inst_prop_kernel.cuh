// inst_prop_kernel.cuh — the fused persistent propagate kernel on the two-level tree (engine_inst.cu),
// built twice like wide_prop_kernel.cuh: engine_inst.cu (`inst_propagate_kernel`) and engine_inst_ool.cu
// (`inst_propagate_ool_kernel`, RNG_DRAW_NOINLINE=1).
#pragma once
#ifndef IPROP_KERNEL_NAME
#define IPROP_KERNEL_NAME inst_propagate_kernel
#endif
// global triangle id of the flattened mesh -> (solid, triangle slot in its template); false if out of range
__device__ __forceinline__ bool inst_locate(const InstView &v, int tri, int &solid, int &slot) {
    if (tri < 0 || (uint32_t)tri >= v.ntriangles) return false;
    const uint32_t s = inst_solid_of_triangle(v.tri_base, v.ninst, (uint32_t)tri);
    solid = (int)s;
    slot = __ldg(v.tri_slot + __ldg(v.tmpl_off_of_orig + s) + ((uint32_t)tri - __ldg(v.tri_base + s)));
    return true;
}

#define IPROP_THREADS 256
#ifndef IPROP_MIN_CTAS
#define IPROP_MIN_CTAS 4
#endif

template <bool STAGED>
__global__ void __launch_bounds__(IPROP_THREADS, IPROP_MIN_CTAS)
IPROP_KERNEL_NAME(GeomView g, InstView v, WideView w, PhotonView ph, WPropParams prm) {
    extern __shared__ float4 smem_raw[];
    uint2 *s_stack = (uint2 *)smem_raw;
    float *s_side = (float *)(s_stack + WIDE_SSTACK * IPROP_THREADS);
    const float *s_opt = nullptr;   // optics tables are read through L1 (engine_core.cu: smem_floats = 0)

    uint2 l_stack[WIDE_STACK - WIDE_SSTACK];
    const WStackRef stk = wstk_make(s_stack + threadIdx.x, l_stack, IPROP_THREADS);
    ISide sd;
    sd.p = s_side + threadIdx.x; sd.nthreads = IPROP_THREADS;
    WSpread spr;
    {
        unsigned char *sb = (unsigned char *)(s_side + ISIDE_WORDS * IPROP_THREADS) + (threadIdx.x >> 5) * WSPREAD_BYTES_PER_WARP;
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
    ITrav tv;
    tv.cur = WIDE_NONE; tv.sp = -1; tv.hit = -1; tv.hit_inst = -1; tv.inst = -1; tv.dist = 0.f;
    uint64_t idx = 0;
    int steps = 0, sfirst = 0;

    while (true) {
        const bool work = have && !itrav_done(tv);
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
                s.distance = tv.hit >= 0 ? tv.dist : -1.0f;
                sd.i(6) = tv.hit;
                p.last_slot = tv.hit;                  // reset by the physics when the photon interacts in the bulk
                if (tv.hit < 0) {                      // photon.h:90-94
                    p.last_tri = -1;
                    p.history |= CHROMA_NO_HIT;
                    done = true;
                } else {
                    const float4 *rec = v.inst + (size_t)tv.hit_inst * 4;
                    const float4 r0 = __ldg(rec), r1 = __ldg(rec + 1), r2 = __ldg(rec + 2), r3 = __ldg(rec + 3);
                    const float4 *tp = v.tris + (size_t)tv.hit * 3;
                    const float4 a = __ldg(tp), b = __ldg(tp + 1), c = __ldg(tp + 2);
                    p.last_tri = (int)(__float_as_uint(r3.z) + __float_as_uint(c.z));
                    sd.i(7) = (int)__float_as_uint(r3.w);          // solid of the triangle just hit (mesh.h:82)
                    const float3 v0 = f3(a.x, a.y, a.z), v1 = f3(a.w, b.x, b.y), v2 = f3(b.z, b.w, c.x);
                    float3 nl = cross3(v1 - v0, v2 - v1);
                    nl = nl / norm3(nl);
                    const float3 nw = f3(r0.x * nl.x + r0.y * nl.y + r0.z * nl.z, r1.x * nl.x + r1.y * nl.y + r1.z * nl.z,
                                         r2.x * nl.x + r2.y * nl.y + r2.z * nl.z);
                    fill_state_optics(s, p, g.opt, s_opt, __float_as_uint(c.y), nw);
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
                    if (p.last_slot < 0) { sd.i(6) = -1; sd.i(7) = -1; }   // photon.h:240,268,273,293
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
                        unsigned long long last = base + (unsigned long long)nfree;
                        if (last > prm.n) last = prm.n;
                        const unsigned long long need = prm.offset + last - prm.sv_offset;
                        unsigned stalled = 0u;
                        if (lane == 0) {
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
                                ph.st[idx] = st;
                                if ((st.x & TERMINAL_MASK) != 0) { ph.pt[idx] = a; ph.dw[idx] = b; ph.pw[idx] = c; }
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
                                int li = -1, ls = -1;
                                inst_locate(v, p.last_tri, li, ls);
                                sd.i(6) = ls; sd.i(7) = li;
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
                    wide_origin_unsearchable(p.pos, v.guard_lim)) {   // wide.cuh: beyond what the grid can hold
                    p.history |= CHROMA_NO_HIT | (unsigned)CHROMA_NAN_ABORT;
                    atomicAdd(prm.nan_count, 1u);
                    ph.pt[idx] = make_float4(p.pos.x, p.pos.y, p.pos.z, p.time);
                    ph.dw[idx] = make_float4(p.dir.x, p.dir.y, p.dir.z, p.wavelength);
                    ph.pw[idx] = make_float4(p.pol.x, p.pol.y, p.pol.z, p.weight);
                    *reinterpret_cast<uint2 *>(ph.st + idx) = make_uint2(p.history, (uint32_t)p.last_tri);
                    have = false;
                    tv.cur = WIDE_NONE;
                } else {
                    itrav_begin(tv, v, p.pos, p.dir);
                }
            }
            continue;
        }
        // ---------------- search phase ----------------
        for (int it = 0; it < prm.iters; it++)
            itrav_round<false>(tv, have, v, w, p.pos, p.dir, stk, sd, spr, prm.leaf_thresh, overflow, dummy0, dummy1);
    }
    if (overflow) atomicOr(prm.status, 1u);
}

