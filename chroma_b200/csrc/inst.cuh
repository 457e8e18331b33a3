// inst.cuh — two-level ("instanced") wide tree and its search.
//
// The reference flattens every placed solid on the host (chroma/geometry.py:360-367:
// np.inner(vertices, rotation) + displacement per solid) and builds ONE tree over the flattened
// mesh: 29,023 copies of the same 5,856-triangle PMT are 168.6 M triangles and 25 GB in the flat
// engine.  Here a solid that is placed many times is stored once (a "template": triangle records in
// its own frame and its own 8-wide tree, wide.cuh format) and the detector is a top tree whose
// leaves are placements (rotation, displacement, template).  A ray that enters a placement's box is
// rotated into the template frame (lengths are preserved, so distances along the ray stay
// comparable), searches the template tree, and returns to the top tree through a sentinel stack
// entry.  Float results differ from the flat mesh in the last bits (the triangle test runs on
// template-frame coordinates), so this is the STATISTICAL tier only and opt-in
// (GPUGeometry(..., instanced=True)); tests/test_gpu_instanced.py compares it with the flat engine
// and with the reference's propagate.cubin.
//
// Layout
//   groups   one array: top tree (root group 0) followed by the template trees; child indices and
//            leaf slots are absolute.  Top-tree leaf: WIDE_LEAF | placement slot (one placement
//            per leaf record).  Template leaf: WIDE_LEAF | (ntri-1) << 28 | first triangle slot.
//   tris     48-byte records of all templates, Morton order inside each template; vertices are
//            relative to the template's box corner; c = (v2.z, material code, triangle id inside
//            the solid, template id)
//   inst     four float4 per top-tree leaf, in leaf order.  A leaf is a placement, or one PATCH of a
//            placement whose box dwarfs its neighbours' (the detector's outer sphere next to 29,023
//            PMTs): such a solid is entered through the subtrees of its template tree a few levels
//            down, each with its own tight box, so that a ray meets the patches in order of distance
//            with everything else instead of searching the whole solid first (its box contains
//            every ray's origin).  Records of the patches differ only in the root group:
//            r0..r2 = rows of the rotation R with the displacement d' in .w
//                     (world = R v + d', d' = displacement + R * template box corner)
//            r3     = (template grid scale, root group, first global triangle id, solid id) bits
#pragma once
#include "wide.cuh"

#ifndef INST_SAVE_FRAME
#define INST_SAVE_FRAME 0   // 1: world-frame slab constants saved on the stack across a placement visit instead of being
                            // recomputed on exit; measured slower (call R, same box: C3 24.39 against 23.63 ms, C4 22.09
                            // against 21.97 ms): three more stack stores per entry and 20 B more spills cost more than
                            // the three reciprocals at 3 - 5 lanes they save
#endif
#define WIDE_EXIT 0xFFFFFFFDu   // stack sentinel: back to the top tree

// InstView: common.cuh

// per-thread side state kept in shared memory (read only in leaf rounds): template-frame ray and the
// excluded triangle; word w of thread t sits at side[w * nthreads + t]
#define ISIDE_WORDS 9   // 0-5 template-frame ray, 6 excluded triangle slot, 7 its solid, 8 solid being searched
struct ISide {
    float *p;
    unsigned nthreads;
    __device__ __forceinline__ float &f(int w) const { return p[w * nthreads]; }
    __device__ __forceinline__ int &i(int w) const { return reinterpret_cast<int *>(p)[w * nthreads]; }
};

struct ITrav {
    float3 A, B;
    unsigned oct;
    unsigned cur;
    int sp;
    float dist;
    int hit;        // best triangle slot (-1 none)
    int hit_inst;   // placement slot of the best hit
    int inst;       // placement being searched (-1: top tree)
};

__device__ __forceinline__ bool itrav_done(const ITrav &t) { return t.cur == WIDE_NONE; }
__device__ __forceinline__ bool itrav_leaf_work(const ITrav &t) {
    return t.cur != WIDE_NONE && (t.cur & WIDE_LEAF) && t.inst >= 0;
}

// slab-test coefficients of a frame whose grid is (origin, scale): wide.cuh wtrav_begin
__device__ __forceinline__ void itrav_frame(ITrav &t, const float3 &o, const float3 &d, const float3 &origin,
                                            float scale) {
    const float ix = wide_inv(d.x), iy = wide_inv(d.y), iz = wide_inv(d.z);
    t.A = f3(scale * ix, scale * iy, scale * iz);
    t.B = f3(__fmaf_rn(-8388608.0f, t.A.x, (origin.x - o.x) * ix),
             __fmaf_rn(-8388608.0f, t.A.y, (origin.y - o.y) * iy),
             __fmaf_rn(-8388608.0f, t.A.z, (origin.z - o.z) * iz));
    t.oct = (__float_as_uint(d.x) >> 31) | (__float_as_uint(d.y) >> 31) << 1 | (__float_as_uint(d.z) >> 31) << 2;
}

__device__ __forceinline__ void itrav_begin(ITrav &t, const InstView &v, const float3 &o, const float3 &d) {
    itrav_frame(t, o, d, v.top_origin, v.top_scale);
    t.sp = -1;
    t.dist = __int_as_float(0x7f800000);
    t.hit = -1;
    t.hit_inst = -1;
    t.inst = -1;
    t.cur = v.ngroups ? v.top_root : WIDE_NONE;
}

__device__ __forceinline__ void itrav_pop(ITrav &t, const WStackRef &k, const InstView &v, const float3 &o,
                                          const float3 &d) {
    t.cur = WIDE_NONE;
    while (t.sp >= 0) {
        uint2 e = wstk_load(k, t.sp--);
        if (e.x == WIDE_EXIT) {          // the placement is exhausted: world frame again
#if INST_SAVE_FRAME
            // the world-frame slab constants were put under the sentinel on entry: three pops instead of
            // recomputing them (three reciprocals, r2s capture: frame set-ups were 9.6 % of all warp
            // instructions at 3 - 5 lanes); the octant follows from the signs of d
            const uint2 e2 = wstk_load(k, t.sp--), e1 = wstk_load(k, t.sp--), e0 = wstk_load(k, t.sp--);
            t.A = f3(__uint_as_float(e0.x), __uint_as_float(e0.y), __uint_as_float(e1.x));
            t.B = f3(__uint_as_float(e1.y), __uint_as_float(e2.x), __uint_as_float(e2.y));
            t.oct = (__float_as_uint(d.x) >> 31) | (__float_as_uint(d.y) >> 31) << 1 | (__float_as_uint(d.z) >> 31) << 2;
#else
            itrav_frame(t, o, d, v.top_origin, v.top_scale);
#endif
            t.inst = -1;
            continue;
        }
        if (!(__uint_as_float(e.y) > t.dist) && (WIDE_POP_FILTER == 0 || e.x != WIDE_EMPTY)) {
            t.cur = e.x;
            break;
        }
    }
}

// top-tree leaf: rotate the ray into the placement's template frame and continue at the template root
// returns false when the stack is full: the placement is skipped and the caller pops (the call fails: overflow)
__device__ __forceinline__ bool itrav_enter(ITrav &t, const InstView &v, const float3 &o, const float3 &d,
                                            const WStackRef &k, const ISide &sd, unsigned &overflow) {
    const unsigned slot = t.cur & 0x0FFFFFFFu;
    const float4 *rec = v.inst + (size_t)slot * 4;
    const float4 r0 = __ldg(rec), r1 = __ldg(rec + 1), r2 = __ldg(rec + 2), r3 = __ldg(rec + 3);
    const float px = o.x - r0.w, py = o.y - r1.w, pz = o.z - r2.w;
    // R^T (o - d'), R^T d
    const float3 ol = f3(r0.x * px + r1.x * py + r2.x * pz, r0.y * px + r1.y * py + r2.y * pz,
                         r0.z * px + r1.z * py + r2.z * pz);
    const float3 dl = f3(r0.x * d.x + r1.x * d.y + r2.x * d.z, r0.y * d.x + r1.y * d.y + r2.y * d.z,
                         r0.z * d.x + r1.z * d.y + r2.z * d.z);
    sd.f(0) = ol.x; sd.f(1) = ol.y; sd.f(2) = ol.z;
    sd.f(3) = dl.x; sd.f(4) = dl.y; sd.f(5) = dl.z;
#if INST_SAVE_FRAME
    if (t.sp + 4 < WIDE_STACK) {     // world-frame A, B under the sentinel (restored by itrav_pop)
        wstk_store(k, t.sp + 1, make_uint2(__float_as_uint(t.A.x), __float_as_uint(t.A.y)));
        wstk_store(k, t.sp + 2, make_uint2(__float_as_uint(t.A.z), __float_as_uint(t.B.x)));
        wstk_store(k, t.sp + 3, make_uint2(__float_as_uint(t.B.y), __float_as_uint(t.B.z)));
        wstk_store(k, t.sp + 4, make_uint2(WIDE_EXIT, 0u));
        t.sp += 4;
    } else {                         // full stack: the call fails (overflow), the search must still terminate
        overflow |= 1u;
        return false;
    }
#else
    const int dep = min(t.sp + 1, WIDE_STACK - 1);
    overflow |= (unsigned)(t.sp + 1 >= WIDE_STACK);
    wstk_store(k, dep, make_uint2(WIDE_EXIT, 0u));
    t.sp = dep;
#endif
    itrav_frame(t, ol, dl, f3(0.f, 0.f, 0.f), r3.x);
    sd.i(8) = (int)__float_as_uint(r3.w);
    t.inst = (int)slot;
    t.cur = __float_as_uint(r3.y);
    return true;
}

// the triangle tests of a template leaf, in the template frame
template <bool COUNT>
__device__ __forceinline__ void itrav_leaf_step(ITrav &t, const InstView &v, const float3 &o, const float3 &d,
                                                const WStackRef &k, const ISide &sd, uint32_t &ntri) {
    const unsigned cur = t.cur;
    const int first = (int)(cur & 0x0FFFFFFFu);
    const int cnt = (int)((cur >> 28) & 7u) + 1;
    RayCtx rc;
    rc.o = f3(sd.f(0), sd.f(1), sd.f(2));
    rc.d = f3(sd.f(3), sd.f(4), sd.f(5));
    // mesh.h:82 last_hit_triangle: a triangle is (solid, slot in its template); a big solid may be in the
    // top tree as several patches, so the solid is compared, not the placement record
    const int skip = (sd.i(7) == sd.i(8)) ? sd.i(6) : -1;
    for (int j = 0; j < cnt; j++) {
        const int slot = first + j;
        const float4 *p = v.tris + (size_t)slot * 3;
        const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
        if (slot != skip) {
            if (COUNT) ntri++;
            float tt;
            if (triangle_hit(rc, f3(a.x, a.y, a.z), f3(a.w, b.x, b.y), f3(b.z, b.w, c.x), tt)) {
                if (tt < t.dist) {
                    t.hit = slot;
                    t.hit_inst = t.inst;
                    t.dist = tt;
                }
            }
        }
    }
    itrav_pop(t, k, v, o, d);
}

// leaf round with the triangle tests dealt over the warp (wide.cuh wtrav_leaf_spread): the template-frame
// rays are already in shared memory (ISide), so a lane reads its pair's ray from the owner's words
template <bool COUNT>
__device__ __forceinline__ void itrav_leaf_spread(ITrav &t, bool lw, const InstView &v, const float3 &o, const float3 &d,
                                                  const WStackRef &k, const ISide &sd, const WSpread &sp, uint32_t &ntri) {
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt = (1u << lane) - 1u;
    const unsigned cnt = lw ? ((t.cur >> 28) & 7u) + 1u : 0u;
    if (__any_sync(0xffffffffu, cnt > 3u)) {
        if (lw) itrav_leaf_step<COUNT>(t, v, o, d, k, sd, ntri);
        return;
    }
    const unsigned b0 = __ballot_sync(0xffffffffu, cnt & 1u), b1 = __ballot_sync(0xffffffffu, cnt & 2u);
    const unsigned off = __popc(b0 & lt) + 2u * __popc(b1 & lt);
    const unsigned total = __popc(b0) + 2u * __popc(b1);
    const unsigned first = t.cur & 0x0FFFFFFFu;
    const int skip = (sd.i(7) == sd.i(8)) ? sd.i(6) : -1;
    for (unsigned j = 0; j < cnt; j++) sp.own[off + j] = (unsigned char)lane;
    __syncwarp();
    for (unsigned base = 0; base < total; base += 32u) {
        const unsigned q = base + lane;
        const bool active = q < total;
        const unsigned src = active ? sp.own[q] : 0u;
        const unsigned o_off = __shfl_sync(0xffffffffu, off, src);
        const unsigned o_first = __shfl_sync(0xffffffffu, first, src);
        const int o_skip = __shfl_sync(0xffffffffu, skip, src);
        if (active) {
            const float *r = sd.p + ((int)src - (int)lane);       // the owner's side words
            RayCtx rc;
            rc.o = f3(r[0], r[sd.nthreads], r[2 * sd.nthreads]);
            rc.d = f3(r[3 * sd.nthreads], r[4 * sd.nthreads], r[5 * sd.nthreads]);
            const int slot = (int)(o_first + (q - o_off));
            float tt = __int_as_float(0x7f800000);
            if (slot != o_skip) {
                const float4 *p = v.tris + (size_t)slot * 3;
                const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
                float th;
                if (triangle_hit(rc, f3(a.x, a.y, a.z), f3(a.w, b.x, b.y), f3(b.z, b.w, c.x), th)) tt = th;
            }
            sp.res[q] = tt;
        }
    }
    __syncwarp();
    if (lw) {
        for (unsigned j = 0; j < cnt; j++) {
            const int slot = (int)(first + j);
            if (COUNT && slot != skip) ntri++;
            const float tt = sp.res[off + j];
            if (tt < t.dist) {
                t.hit = slot;
                t.hit_inst = t.inst;
                t.dist = tt;
            }
        }
        itrav_pop(t, k, v, o, d);
    }
    __syncwarp();
}

// Warp-level search round (wide.cuh wtrav_round) with a third kind of step: ENTERING a placement (a
// top-tree leaf: rotate the ray into the template frame, ~90 instructions).  r2c capture: done inside
// the group round it ran at 3.2 lanes for 7.6 % of all warp instructions; lanes that reached a
// top-tree leaf now wait, like lanes that reached a triangle leaf, until `leaf_thresh` of them are
// ready or nothing else can run.
#ifndef INST_BATCH_ENTER
#define INST_BATCH_ENTER 0   // r2e: batching the enters (27.7 ms) lost to entering inside the group round (24.6 ms)
#endif
#ifndef INST_LEAF_SPREAD
#define INST_LEAF_SPREAD WIDE_LEAF_SPREAD
#endif
template <bool COUNT>
__device__ __forceinline__ void itrav_round(ITrav &t, bool searching, const InstView &v, const WideView &w,
                                            const float3 &o, const float3 &d, const WStackRef &k, const ISide &sd,
                                            const WSpread &sp, int leaf_thresh, unsigned &overflow, uint32_t &ngroup,
                                            uint32_t &ntri) {
    const bool lw = searching && itrav_leaf_work(t);
    const bool ew = searching && t.cur != WIDE_NONE && (t.cur & WIDE_LEAF) && t.inst < 0;
    const bool gw = searching && t.cur != WIDE_NONE && !(t.cur & WIDE_LEAF);
    const unsigned mL = __ballot_sync(0xffffffffu, lw);
    const unsigned mE = __ballot_sync(0xffffffffu, ew);
    const unsigned mG = __ballot_sync(0xffffffffu, gw);
    const int nL = __popc(mL), nE = __popc(mE);
#if INST_BATCH_ENTER
    if (nE >= leaf_thresh || (mG == 0u && nE > 0 && nE >= nL)) {
        if (ew) itrav_enter(t, v, o, d, k, sd, overflow);
    } else
#endif
    if ((INST_BATCH_ENTER ? mG : (mG | mE)) == 0u || nL >= leaf_thresh) {
#if INST_LEAF_SPREAD
        itrav_leaf_spread<COUNT>(t, lw, v, o, d, k, sd, sp, ntri);
#else
        if (lw) itrav_leaf_step<COUNT>(t, v, o, d, k, sd, ntri);
#endif
    } else if (gw || ew) {
#if !INST_BATCH_ENTER
        bool entered = true;
        if (ew) entered = itrav_enter(t, v, o, d, k, sd, overflow);     // then the box tests of the template's root group
        if (!entered) itrav_pop(t, k, v, o, d);
        else
#else
        if (gw)
#endif
        if (wtrav_group_core<COUNT>(t, w, k, overflow, ngroup)) itrav_pop(t, k, v, o, d);
    }
}
