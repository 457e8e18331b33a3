// wide_prop2.cuh — propagate kernel with two photons per lane (CHROMA_PROP_V=5).
//
// Same physics, same per-photon random streams and the same search as wide_propagate_kernel
// (engine_wide.cu), scheduled differently.  There a lane owns one photon, so a lane whose search has
// ended idles until enough lanes of its warp are ready for a physics phase (r02d profile: group
// rounds run with 14.5 of 32 lanes).  Here a lane owns two photon slots: while one photon waits for
// its physics the other one is searched, so search rounds keep (nearly) every lane busy and the
// physics phase still runs over >= `refill` lanes.  Everything a photon carries between its search
// and its physics lives in a per-lane global side buffer (one 128-byte line per slot, ~39 MB);
// only the ray and the search state stay in registers.
#pragma once

#define W2_WORDS 32      // one 128-byte line per slot
#ifndef W2_MIN_CTAS
#define W2_MIN_CTAS 4
#endif
enum { W2_EMPTY = 0, W2_READY = 1, W2_SEARCH = 2, W2_DONE = 3 };

// A slot is 8 x 16 bytes, contiguous per lane (whole sectors whichever lanes of the warp take part):
//   q0 pos, time | q1 dir, wavelength | q2 pol, weight | q3 history, last slot, last triangle, steps
//   q4 rng blk, left, b0, b1 | q5 rng b2, scatter_first, idx lo, idx hi | q6 search result hit, distance
struct W2Side {
    uint4 *base;   // slot s of this lane: base + s * 8
    __device__ __forceinline__ uint4 *slot(int s) const { return base + s * 8; }
    __device__ __forceinline__ void set_result(int s, int hit, float dist) const {
        *reinterpret_cast<uint2 *>(slot(s) + 6) = make_uint2((uint32_t)hit, __float_as_uint(dist));
    }
};
__device__ __forceinline__ uint4 w2_pack(float x, float y, float z, float w) {
    return make_uint4(__float_as_uint(x), __float_as_uint(y), __float_as_uint(z), __float_as_uint(w));
}

__device__ __forceinline__ void w2_store_photon(const W2Side &sd, int s, const Photon &p, const Rng &rng, int steps,
                                                int sfirst, uint64_t idx) {
    uint4 *q = sd.slot(s);
    q[0] = w2_pack(p.pos.x, p.pos.y, p.pos.z, p.time);
    q[1] = w2_pack(p.dir.x, p.dir.y, p.dir.z, p.wavelength);
    q[2] = w2_pack(p.pol.x, p.pol.y, p.pol.z, p.weight);
    q[3] = make_uint4(p.history, (uint32_t)p.last_slot, (uint32_t)p.last_tri, (uint32_t)steps);
    q[4] = make_uint4(rng.blk, rng.left, rng.b0, rng.b1);
    q[5] = make_uint4(rng.b2, (uint32_t)sfirst, (uint32_t)idx, (uint32_t)(idx >> 32));
}

__device__ __forceinline__ void w2_load_photon(const W2Side &sd, int s, Photon &p, Rng &rng, int &steps, int &sfirst,
                                               uint64_t &idx, uint32_t id_base, int &hit, float &dist) {
    const uint4 *q = sd.slot(s);
    const uint4 a = q[0], b = q[1], c = q[2], e = q[3], f = q[4], h = q[5];
    const uint2 r = *reinterpret_cast<const uint2 *>(q + 6);
    p.pos = f3(__uint_as_float(a.x), __uint_as_float(a.y), __uint_as_float(a.z)); p.time = __uint_as_float(a.w);
    p.dir = f3(__uint_as_float(b.x), __uint_as_float(b.y), __uint_as_float(b.z)); p.wavelength = __uint_as_float(b.w);
    p.pol = f3(__uint_as_float(c.x), __uint_as_float(c.y), __uint_as_float(c.z)); p.weight = __uint_as_float(c.w);
    p.history = e.x; p.last_slot = (int)e.y; p.last_tri = (int)e.z; steps = (int)e.w;
    rng.blk = f.x; rng.left = f.y; rng.b0 = f.z; rng.b1 = f.w; rng.b2 = h.x;
    sfirst = (int)h.y;
    idx = (uint64_t)h.z | (uint64_t)h.w << 32;
    rng.id = id_base + (uint32_t)idx;
    hit = (int)r.x; dist = __uint_as_float(r.y);
}

__device__ __forceinline__ void w2_retire(const PhotonView &ph, uint64_t idx, const Photon &p) {
    ph.pt[idx] = make_float4(p.pos.x, p.pos.y, p.pos.z, p.time);
    ph.dw[idx] = make_float4(p.dir.x, p.dir.y, p.dir.z, p.wavelength);
    ph.pw[idx] = make_float4(p.pol.x, p.pol.y, p.pol.z, p.weight);
    *reinterpret_cast<uint2 *>(ph.st + idx) = make_uint2(p.history, (uint32_t)p.last_tri);
}

template <bool STAGED>
__global__ void __launch_bounds__(WPROP_THREADS, W2_MIN_CTAS)
wide_propagate2_kernel(GeomView g, WideView w, PhotonView ph, WPropParams prm, uint32_t *side_base) {
    extern __shared__ float4 smem_raw[];
    uint2 *s_stack = (uint2 *)smem_raw;
    float *s_opt = (float *)(s_stack + WIDE_SSTACK * WPROP_THREADS);
    for (uint32_t i = threadIdx.x; i < g.opt.smem_floats; i += blockDim.x) s_opt[i] = __ldg(g.opt.blob + i);
    __syncthreads();

    uint2 l_stack[WIDE_STACK - WIDE_SSTACK];
    const WStackRef stk = wstk_make(s_stack + threadIdx.x, l_stack, WPROP_THREADS);
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    const bool use_weights = prm.use_weights != 0;
    const RngKey K = rng_key(prm.seed, prm.epoch, 0u);
    W2Side sd;
    sd.base = reinterpret_cast<uint4 *>(side_base) + ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 16;

    int st0 = W2_EMPTY, st1 = W2_EMPTY;   // slot states
    bool exhausted = false;
    unsigned overflow = 0u;
    uint32_t dummy0 = 0, dummy1 = 0;
    float3 o = f3(0.f, 0.f, 0.f), d = f3(0.f, 0.f, 1.f);
    WTrav tv;
    tv.cur = WIDE_NONE; tv.sp = -1; tv.hit = -1; tv.dist = 0.f; tv.last_slot = -1;
    tv.A = o; tv.B = o; tv.oct = 0u;
    // ray of the slot that is READY (at most one at a time: a ready slot is taken as soon as the
    // lane's search ends), kept in registers so that switching photons costs no memory round trip
    float3 ro = o, rd = d;
    int rls = -1;

    // a finished search parks its result; a lane without a search takes its ready slot
    auto lane_switch = [&]() {
        if (tv.cur == WIDE_NONE) {
            if (st0 == W2_SEARCH || st1 == W2_SEARCH) {
                const int c = st0 == W2_SEARCH ? 0 : 1;
                sd.set_result(c, tv.hit, tv.dist);
                if (c == 0) st0 = W2_DONE; else st1 = W2_DONE;
            }
            if (st0 == W2_READY || st1 == W2_READY) {
                const int r = st0 == W2_READY ? 0 : 1;
                o = ro;
                d = rd;
                wtrav_begin(tv, w, o, d, rls);
                if (wide_origin_unsearchable(o, WIDE_GUARD_QUANTA * w.world_scale)) tv.cur = WIDE_NONE;   // wide.cuh
                if (r == 0) st0 = W2_SEARCH; else st1 = W2_SEARCH;
                if (tv.cur == WIDE_NONE) {       // empty tree: the search is over at once
                    sd.set_result(r, tv.hit, tv.dist);
                    if (r == 0) st0 = W2_DONE; else st1 = W2_DONE;
                }
            }
        }
    };

    while (true) {
        lane_switch();
        const bool searching = tv.cur != WIDE_NONE;
        const bool pending = st0 == W2_DONE || st1 == W2_DONE ||
                             (!exhausted && (st0 == W2_EMPTY || st1 == W2_EMPTY));
        const unsigned mS = __ballot_sync(0xffffffffu, searching);
        const unsigned mP = __ballot_sync(0xffffffffu, pending);
        const unsigned mA = __ballot_sync(0xffffffffu, st0 != W2_EMPTY || st1 != W2_EMPTY);
        if (mA == 0u && exhausted) break;

        if (mS == 0u || __popc(mP) >= prm.refill) {
            // ---------------- physics phase: one slot per lane ----------------
            int c = st0 == W2_DONE ? 0 : (st1 == W2_DONE ? 1 : -1);
            int rslot = -1;
            if (c < 0 && !exhausted) rslot = st0 == W2_EMPTY ? 0 : (st1 == W2_EMPTY ? 1 : -1);
            Photon p;
            Rng rng;
            int steps = 0, sfirst = 0;
            uint64_t idx = 0;
            bool start = false;
            if (c >= 0) {
                int hit;
                float dist;
                w2_load_photon(sd, c, p, rng, steps, sfirst, idx, prm.id_base, hit, dist);
                bool done;
                State s;
                if (!fill_state_post(s, p, g, s_opt, hit, hit >= 0 ? dist : -1.0f)) {
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
                    w2_retire(ph, idx, p);
                    if (c == 0) st0 = W2_EMPTY; else st1 = W2_EMPTY;
                    if (!exhausted) rslot = c;
                } else {
                    start = true;
                }
            }
            if (!exhausted) {
                const bool want = rslot >= 0;
                const unsigned wm = __ballot_sync(0xffffffffu, want);
                const int nfree = __popc(wm);
                if (nfree > 0) {
                    unsigned long long base = 0;
                    if (lane == 0) base = atomicAdd(prm.head, (unsigned long long)nfree);
                    base = __shfl_sync(0xffffffffu, base, 0);
                    if (base + (unsigned long long)nfree >= prm.n) exhausted = true;
                    if (STAGED && base < prm.n) {
                        unsigned long long last = base + (unsigned long long)nfree;
                        if (last > prm.n) last = prm.n;
                        const unsigned long long need = prm.offset + last - prm.sv_offset;
                        if (lane == 0) {
                            unsigned spins = 0;   // watchdog, as in wide_propagate_kernel
                            while (*(volatile const unsigned long long *)prm.ready < need) {
                                __nanosleep(1000);
                                if (++spins > 4000000u) { atomicOr(prm.status, 2u); break; }
                            }
                        }
                        __syncwarp();
                    }
                    if (want) {
                        unsigned long long i = base + (unsigned long long)__popc(wm & lt_mask);
                        if (i < prm.n) {
                            idx = prm.offset + i;
                            uint4 st;
                            float4 a, b, cc;
                            if (STAGED) {
                                const StageView &sv = prm.sv;
                                const uint64_t k = idx - prm.sv_offset;
                                st = make_uint4(sv.flags ? __ldcg(sv.flags + k) : 0u,
                                                (uint32_t)(sv.last_hit ? __ldcg(sv.last_hit + k) : -1),
                                                sv.evidx ? __ldcg(sv.evidx + k) : 0u, 0xFFFFFFFFu);
                                a = make_float4(__ldcg(sv.pos + 3 * k), __ldcg(sv.pos + 3 * k + 1), __ldcg(sv.pos + 3 * k + 2), __ldcg(sv.t + k));
                                b = make_float4(__ldcg(sv.dir + 3 * k), __ldcg(sv.dir + 3 * k + 1), __ldcg(sv.dir + 3 * k + 2), __ldcg(sv.wl + k));
                                cc = make_float4(__ldcg(sv.pol + 3 * k), __ldcg(sv.pol + 3 * k + 1), __ldcg(sv.pol + 3 * k + 2),
                                                 sv.weights ? __ldcg(sv.weights + k) : 1.0f);
                                ph.st[idx] = st;
                                if ((st.x & TERMINAL_MASK) != 0) { ph.pt[idx] = a; ph.dw[idx] = b; ph.pw[idx] = cc; }
                            } else {
                                st = ph.st[idx];
                                if ((st.x & TERMINAL_MASK) == 0) { a = ph.pt[idx]; b = ph.dw[idx]; cc = ph.pw[idx]; }
                            }
                            if ((st.x & TERMINAL_MASK) == 0) {  // propagate.cu:258
                                p.pos = f3(a.x, a.y, a.z); p.time = a.w;
                                p.dir = f3(b.x, b.y, b.z); p.wavelength = b.w;
                                p.pol = f3(cc.x, cc.y, cc.z); p.weight = cc.w;
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
                                c = rslot;
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
                    w2_retire(ph, idx, p);
                    if (c == 0) st0 = W2_EMPTY; else st1 = W2_EMPTY;
                } else {
                    w2_store_photon(sd, c, p, rng, steps, sfirst, idx);
                    if (c == 0) st0 = W2_READY; else st1 = W2_READY;
                    ro = p.pos; rd = p.dir; rls = p.last_slot;
                }
            }
            continue;
        }
        // ---------------- search phase ----------------
        for (int it = 0; it < prm.iters; it++) {
            wtrav_round<false>(tv, true, w, o, d, stk, prm.leaf_thresh, overflow, dummy0, dummy1, WSpread{nullptr, nullptr});
            lane_switch();
        }
    }
    if (overflow) atomicOr(prm.status, 1u);
}
