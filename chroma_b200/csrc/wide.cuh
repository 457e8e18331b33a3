// wide.cuh — the engine's own acceleration structure ("wide tree") and its near-first search.
//
// Why a second tree.  The reference searches its Morton-grid BVH (chroma/bvh/grid.py:11-95,
// degree ~3, up to 15 children, one triangle per leaf) depth first in storage order
// (chroma/cuda/mesh.h:41-118): ~270 dependent 16-byte node fetches per ray in the 29k-PMT
// detector, most of them behind the hit.  What the reference RETURNS, however, does not depend
// on that order.  Its answer is
//        argmin over triangles accepted by intersect_triangle() of (distance, search rank)
// where "search rank" is the position of the triangle's leaf in the reference's fixed visiting
// order (leaves of a sibling group in index order, then the inner siblings last-to-first, each
// with its whole subtree): a later triangle replaces the current one only if `distance <
// min_distance` (mesh.h:88), and a box is skipped only if it cannot contain a nearer hit
// (mesh.h:26-30; the 1-quantum padding of every leaf box, bvh.cu:186-195, is ~100x larger than
// the rounding error of the slab test, see DESIGN.md "Search equivalence").
// So the engine is free to search any conservative hierarchy in any order as long as
//   (1) the triangle test is the reference's, bit for bit (traverse.cuh triangle_hit), and
//   (2) ties in distance are resolved by the reference's search rank (stored per leaf slot).
// tests/test_reference_gpu.py checks the result against the reference's own kernels.
//
// Layout.  A group is eight 16-byte records in one 128-byte line (one L2 line / four sectors):
//   record = uint4 { x, y, z : lo | hi << 16 on the reference's 16-bit world grid
//                              (geometry.h:30-47, same decode),
//                    w       : inner  -> index of the child group          (bit 31 clear)
//                              leaf   -> 1<<31 | (ntri-1) << 28 | first slot
//                              unused -> 0xFFFFFFFF (unused records are at the end) }
// Triangles of a leaf are consecutive 48-byte records in Morton order.  Groups are built top
// down over the sorted Morton codes of the triangle centroids (engine_wide.cu).
#pragma once
#include "traverse.cuh"

#define WIDE_EMPTY 0xFFFFFFFFu
#define WIDE_LEAF 0x80000000u
#define WIDE_NONE 0xFFFFFFFEu  // "no current node" marker of the search state

#ifndef WIDE_SSTACK
#define WIDE_SSTACK 20          // stack entries per thread kept in shared memory (r01v: 12 -> 20 = +6 %,
#endif                          // spilled entries missed L1 83 % of the time)
#define WIDE_STACK 96
#ifndef WIDE_LEAF_BATCH
#define WIDE_LEAF_BATCH 1       // triangle records of a leaf requested together (r01z: 2 or 3 spill at 64 registers and lose 40-65 %)
#endif           // total entries (rest spills to local memory)

// Prefetch on accept: every child whose box the ray enters is requested right away (its group
// line, or the triangle records of a leaf), so the misses of one group's children overlap instead
// of being taken one after the other when each child is popped (r01t profile: 40 % of the stall
// samples were first uses of a group line or a triangle record).  0 none, 1 L2, 2 L1.
#ifndef WIDE_PRED_PUSH
#define WIDE_PRED_PUSH 0   // r02f: branch-free push with a predicated store = same time (278 vs 278 Mphot/s), kept for A/B
#endif
#ifndef WIDE_EMPTY_CHECK
// 1: test r.w != WIDE_EMPTY in the two-pass group step.  The test is redundant for every ray the grid can hold (an
// empty record carries the inverted box) and dropping it gave 1.5 % (call K), but then a ray that is NOT in that
// class (garbage read after a stalled upload, call Z) followed an empty record's index out of the arrays; the
// filters that make the untested form memory-safe whatever the ray (WIDE_POP_FILTER) cost more than the test
// itself on the flat kernel (same box, call T: 26.35 ms against 25.30 ms; two-level 23.50 against 23.61).
#define WIDE_EMPTY_CHECK 1
#endif
// with the empty-record test off, an empty record's index is filtered where the nearest child is selected and
// at pop (memory safety whatever the ray); with the test on it never gets that far
#define WIDE_POP_FILTER (WIDE_TWO_PASS && !WIDE_EMPTY_CHECK)
#ifndef WIDE_FAST_PUSH
#define WIDE_FAST_PUSH 1   // predicated pushes of the two-pass group step while the stack is in shared memory (r2j)
#endif
#ifndef WIDE_TWO_PASS
#define WIDE_TWO_PASS 1   // see wtrav_group_core; r2b: alone 27.57 vs 27.15 ms, with WIDE_LEAF_SPREAD 26.40 vs 26.64
#endif
#ifndef WIDE_LD256
#define WIDE_LD256 1
#endif
#ifndef WIDE_PF
#define WIDE_PF 0   // r01u: no gain on B200 (the misses are not what serialises), kept for A/B
#endif
__device__ __forceinline__ void wide_prefetch(const void *p) {
#if WIDE_PF == 1
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#elif WIDE_PF == 2
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#endif
}
__device__ __forceinline__ void wide_prefetch_child(const WideView &w, unsigned ref) {
#if WIDE_PF
    if (ref & WIDE_LEAF) {
        const char *p = (const char *)(w.tris + (size_t)(ref & 0x0FFFFFFFu) * 3);
        wide_prefetch(p);
        wide_prefetch(p + 48 * (((ref >> 28) & 7u) + 1u) - 16);   // last 16 bytes of the leaf's records
    } else {
        wide_prefetch(w.groups + (size_t)ref * 8);
    }
#endif
}

struct WStackRef {
    uint2 *sh;       // shared: sh[depth * nthreads] (already offset by the thread index)
    uint2 *lo;       // local spill: lo[depth - WIDE_SSTACK]
    unsigned nthreads;
    unsigned sh32;   // 32-bit shared-window address of sh (wstk_make)
};
// The address comes from a volatile asm so that it is computed once and kept: left to the compiler it is
// rematerialised at every push and pop (S2R SR_TID.X + S2R SR_CgaCtaId + three moves, s19x capture).
__device__ __forceinline__ WStackRef wstk_make(uint2 *sh, uint2 *lo, unsigned nthreads) {
    WStackRef k;
    k.sh = sh; k.lo = lo; k.nthreads = nthreads;
    asm volatile("{\n\t.reg .u64 t;\n\tcvta.to.shared.u64 t, %1;\n\tcvt.u32.u64 %0, t;\n\t}" : "=r"(k.sh32) : "l"(sh));
    return k;
}
// Stack access.  WIDE_STS_ASM=1 (default): entries below WIDE_SSTACK go through explicit st.shared /
// ld.shared with a 32-bit shared-window address.  The generic-pointer form (0) made every push
// convert the shared address to a generic one (three S2R: SR_TID.X, SR_CgaCtaId, SR_SWINHI, and a
// dozen moves), and the push runs at 3-4 lanes for 22 % of all warp instructions (s19y capture).
#ifndef WIDE_STS_ASM
#define WIDE_STS_ASM 1
#endif
__device__ __forceinline__ uint2 *wstk_ptr(const WStackRef &k, int depth) {
    return depth < WIDE_SSTACK ? k.sh + depth * k.nthreads : k.lo + (depth - WIDE_SSTACK);
}
__device__ __forceinline__ void wstk_store(const WStackRef &k, int depth, uint2 e) {
#if WIDE_STS_ASM
    if (depth < WIDE_SSTACK) {
        const unsigned a = k.sh32 + (unsigned)depth * (k.nthreads * 8u);
        asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(a), "r"(e.x), "r"(e.y) : "memory");
    } else {
        k.lo[depth - WIDE_SSTACK] = e;
    }
#else
    *wstk_ptr(k, depth) = e;
#endif
}
__device__ __forceinline__ uint2 wstk_load(const WStackRef &k, int depth) {
#if WIDE_STS_ASM
    uint2 e;
    if (depth < WIDE_SSTACK) {
        const unsigned a = k.sh32 + (unsigned)depth * (k.nthreads * 8u);
        asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(e.x), "=r"(e.y) : "r"(a) : "memory");
    } else {
        e = k.lo[depth - WIDE_SSTACK];
    }
    return e;
#else
    return *wstk_ptr(k, depth);
#endif
}

struct WTrav {
    float3 A, B;      // slab test: t = (2^23 + u16) * A + B  (A = scale/d, B = (origin - o)/d - 2^23 A)
    unsigned oct;     // bit k set: d.k < 0 (the upper bound is the near one on that axis)
    unsigned cur;     // record to process next (group index, leaf word) or WIDE_NONE
    int sp;           // stack top (-1 empty)
    float dist;       // best distance (+inf none)
    int hit;          // best slot (-1 none)
    int last_slot;    // slot excluded from the search (mesh.h:82)
};

__device__ __forceinline__ bool wtrav_done(const WTrav &t) { return t.cur == WIDE_NONE; }

// 1/d for the slab test; |d| below 1e-18 is treated as 1e-18 so that rays parallel to an axis
// keep a finite, huge slope: boxes that do not contain the origin's coordinate are rejected,
// boxes that do put no constraint
__device__ __forceinline__ float wide_inv(float d) {
    float a = fabsf(d);
    float r = __fdividef(1.0f, fmaxf(a, 1e-18f));
    return copysignf(r, d);
}

// A ray the 16-bit grid cannot hold: origin farther than ~2^34 quanta from the grid (or not finite).  For every
// other ray an INVERTED box (lo = 65535, hi = 0: the box of an empty record) fails the slab test on each axis
// by itself, because near - far = 65535 |A| cannot vanish against |B| <= (|origin - o| / scale + 2^23) |A| in a
// float (it would take 2^39 quanta); that is the argument behind WIDE_EMPTY_CHECK = 0 (not the default, see there).
// Directions are unit vectors or NaN at every call site (normalised on load), and a NaN component only removes its
// axis from the test.  The kernels test the origin where a ray is loaded / a step starts (next to
// propagate.cu:270-273's NaN test) and end such a ray as "no hit": a photon uploaded with infinite or absurd
// coordinates would otherwise pass every box (near and far planes round to the same value) and walk the whole tree.
#define WIDE_GUARD_QUANTA 1.7e10f   // 2^34
__device__ __forceinline__ bool wide_origin_unsearchable(const float3 &o, float lim) {
    return !(fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z)) < lim);
}

__device__ __forceinline__ void wtrav_begin(WTrav &t, const WideView &w, const float3 &o,
                                            const float3 &d, int last_slot) {
    const float ix = wide_inv(d.x), iy = wide_inv(d.y), iz = wide_inv(d.z);
    t.A = f3(w.world_scale * ix, w.world_scale * iy, w.world_scale * iz);
    // The 16-bit bounds are turned into floats by byte-permuting them under the exponent of 2^23
    // (0x4B000000 | u = 8388608 + u exactly); the bias is folded into B with ONE rounding
    // (<= 0.5 |A|, i.e. half a grid quantum in t) so the test is a single FMA per bound.  The wide
    // tree pads its boxes by 2 quanta (engine_wide.cu quant_lo/hi) to keep the slab test conservative.
    t.B = f3(__fmaf_rn(-8388608.0f, t.A.x, (w.world_origin.x - o.x) * ix),
             __fmaf_rn(-8388608.0f, t.A.y, (w.world_origin.y - o.y) * iy),
             __fmaf_rn(-8388608.0f, t.A.z, (w.world_origin.z - o.z) * iz));
    t.oct = (__float_as_uint(d.x) >> 31) | (__float_as_uint(d.y) >> 31) << 1 | (__float_as_uint(d.z) >> 31) << 2;
    t.sp = -1;
    t.dist = __int_as_float(0x7f800000);
    t.hit = -1;
    t.last_slot = last_slot;
    t.cur = w.ngroups ? 0u : WIDE_NONE;   // group 0 is the root group
}

// next record: the nearest child just found, else the nearest-first stack (entries farther than
// the best hit are dropped)
__device__ __forceinline__ void wtrav_pop(WTrav &t, const WStackRef &k) {
    t.cur = WIDE_NONE;
    while (t.sp >= 0) {
        uint2 e = wstk_load(k, t.sp--);
        if (!(__uint_as_float(e.y) > t.dist) && (WIDE_POP_FILTER == 0 || e.x != WIDE_EMPTY)) {
            t.cur = e.x;
            break;
        }
    }
}

// what a lane needs next
__device__ __forceinline__ bool wtrav_leaf_work(const WTrav &t) { return t.cur != WIDE_NONE && (t.cur & WIDE_LEAF); }
__device__ __forceinline__ bool wtrav_group_work(const WTrav &t) { return !(t.cur & WIDE_LEAF); }

// leaf micro-step: the triangle tests of the current leaf (caller guarantees wtrav_leaf_work).
// A whole leaf is resolved in one round (r01z: +4 % over one triangle per round), so a leaf costs one memory round
// trip instead of one per triangle (r01t profile: the first use of a triangle record was 19 % of
// all stall samples).
template <bool COUNT>
__device__ __forceinline__ void wtrav_leaf_step(WTrav &t, const WideView &w, const float3 &o, const float3 &d,
                                                const WStackRef &k, uint32_t &ntri) {
    const unsigned cur = t.cur;
    const int first = (int)(cur & 0x0FFFFFFFu);
    const int cnt = (int)((cur >> 28) & 7u) + 1;
    RayCtx rc;
    rc.o = o;
    rc.d = d;
    for (int base = 0; base < cnt; base += WIDE_LEAF_BATCH) {
        float4 a[WIDE_LEAF_BATCH], b[WIDE_LEAF_BATCH], c[WIDE_LEAF_BATCH];
#pragma unroll
        for (int j = 0; j < WIDE_LEAF_BATCH; j++) {
            if (base + j < cnt) {
                const float4 *p = w.tris + (size_t)(first + base + j) * 3;
                a[j] = __ldg(p); b[j] = __ldg(p + 1); c[j] = __ldg(p + 2);
            }
        }
#pragma unroll
        for (int j = 0; j < WIDE_LEAF_BATCH; j++) {
            const int slot = first + base + j;
            if (base + j < cnt && slot != t.last_slot) {
                if (COUNT) ntri++;
                float tt;
                if (triangle_hit(rc, f3(a[j].x, a[j].y, a[j].z), f3(a[j].w, b[j].x, b[j].y),
                                 f3(b[j].z, b[j].w, c[j].x), tt)) {
                    // mesh.h:88 keeps the first of equal distances in ITS visiting order: lower rank wins
                    if (tt < t.dist || (tt == t.dist && t.hit >= 0 && __ldg(w.rank + slot) < __ldg(w.rank + t.hit))) {
                        t.hit = slot;
                        t.dist = tt;
                    }
                }
            }
        }
    }
    wtrav_pop(t, k);
}

// group micro-step: the 8 box tests of the current group (caller guarantees wtrav_group_work).
// Generic over the search state (WTrav here, ITrav in inst.cuh): returns true when no child was
// entered and the caller has to pop.
template <bool COUNT, class TRAV>
__device__ __forceinline__ bool wtrav_group_core(TRAV &t, const WideView &w, const WStackRef &k,
                                                 unsigned &overflow, uint32_t &ngroup) {
    if (COUNT) ngroup++;
    const uint4 *grp = w.groups + (size_t)t.cur * 8;
    uint4 r[8];
#if WIDE_LD256
    // four 256-bit loads (LDG.E.256, sm_100a) instead of eight 128-bit ones: half the load
    // instructions and L1 wavefronts for the same 128-byte line
#pragma unroll
    for (int j = 0; j < 4; j++)
        asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(r[2 * j].x), "=r"(r[2 * j].y), "=r"(r[2 * j].z), "=r"(r[2 * j].w),
                       "=r"(r[2 * j + 1].x), "=r"(r[2 * j + 1].y), "=r"(r[2 * j + 1].z), "=r"(r[2 * j + 1].w)
                     : "l"(grp + 2 * j));
#else
#pragma unroll
    for (int j = 0; j < 8; j++) r[j] = __ldg(grp + j);
#endif
    float best_t = __int_as_float(0x7f800000);
    unsigned best = WIDE_NONE;
    // near / far bound of each axis picked by the ray's octant: no per-axis min/max
    const unsigned nx = (t.oct & 1u) ? 0x7632u : 0x7610u, fx = nx ^ 0x0022u;
    const unsigned ny = (t.oct & 2u) ? 0x7632u : 0x7610u, fy = ny ^ 0x0022u;
    const unsigned nz = (t.oct & 4u) ? 0x7632u : 0x7610u, fz = nz ^ 0x0022u;
#if WIDE_TWO_PASS
    // pass 1 runs the eight slab tests branch-free into a hit mask, pass 2 picks the nearest child and
    // pushes the others (round 2: default on, measured together with the leaf spread).  Motivation: BRA + BSYNC + BSSY were 17 % of the warp instructions of the final
    // capture, one divergent block per child.  The order of the pushes differs from the streaming form
    // (child order, nearest excluded); the result does not depend on it.
    float tnv[8];
    unsigned hitmask = 0u;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const float tnx = __fmaf_rn(__uint_as_float(__byte_perm(r[j].x, 0x4B000000u, nx)), t.A.x, t.B.x);
        const float tfx = __fmaf_rn(__uint_as_float(__byte_perm(r[j].x, 0x4B000000u, fx)), t.A.x, t.B.x);
        const float tny = __fmaf_rn(__uint_as_float(__byte_perm(r[j].y, 0x4B000000u, ny)), t.A.y, t.B.y);
        const float tfy = __fmaf_rn(__uint_as_float(__byte_perm(r[j].y, 0x4B000000u, fy)), t.A.y, t.B.y);
        const float tnz = __fmaf_rn(__uint_as_float(__byte_perm(r[j].z, 0x4B000000u, nz)), t.A.z, t.B.z);
        const float tfz = __fmaf_rn(__uint_as_float(__byte_perm(r[j].z, 0x4B000000u, fz)), t.A.z, t.B.z);
        const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));
        const float tf = fminf(fminf(tfx, tfy), tfz);
        tnv[j] = tn;
#if WIDE_EMPTY_CHECK
        hitmask |= (r[j].w != WIDE_EMPTY && tn <= tf && !(tn > t.dist)) ? (1u << j) : 0u;
#else
        // an empty record carries the inverted box (lo = 65535, hi = 0 on every axis: wk_level_kernel), so
        // tn > tf for any ray and the record needs no test of its own
        hitmask |= (tn <= tf && !(tn > t.dist)) ? (1u << j) : 0u;
#endif
    }
    int bj = -1;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const bool c = ((hitmask >> j) & 1u) && tnv[j] < best_t;
        best_t = c ? tnv[j] : best_t;
        best = c ? r[j].w : best;
        bj = c ? j : bj;
    }
    if (bj >= 0) hitmask &= ~(1u << bj);
#if WIDE_TWO_PASS == 2
    // pushes without a branch: one predicated shared store per child (entries beyond the shared part of
    // the stack take the slow path)
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const unsigned push = (hitmask >> j) & 1u;
        const int d = min(t.sp + 1, WIDE_STACK - 1);
        overflow |= (unsigned)(push && t.sp + 1 >= WIDE_STACK);
        const unsigned sa = k.sh32 + (unsigned)d * (k.nthreads * 8u);
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %0, 0;\n\t@p st.shared.v2.u32 [%1], {%2, %3};\n\t}"
                     :: "r"((unsigned)(push && d < WIDE_SSTACK)), "r"(sa), "r"(r[j].w), "r"(__float_as_uint(tnv[j])) : "memory");
        if (push && d >= WIDE_SSTACK) k.lo[d - WIDE_SSTACK] = make_uint2(r[j].w, __float_as_uint(tnv[j]));
        t.sp = push ? d : t.sp;
    }
#else
#if WIDE_FAST_PUSH
    // r2i capture: the per-child push blocks (address, bounds, store, each under its own divergent branch)
    // were 16 % of all warp instructions at 3-4 lanes.  When all pushes of this group land in the
    // shared part of the stack (the common case, one test) they become three predicated instructions
    // per child, run by all lanes together: test bit j, store at the running address, advance it.
    const int npush = __popc(hitmask);
    if (t.sp + npush < WIDE_SSTACK) {
        unsigned sa = k.sh32 + (unsigned)(t.sp + 1) * (k.nthreads * 8u);
#pragma unroll
        for (int j = 0; j < 8; j++)
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %1, 0;\n\t@p st.shared.v2.u32 [%0], {%2, %3};\n\t@p add.u32 %0, %0, %4;\n\t}"
                         : "+r"(sa)
                         : "r"(hitmask & (1u << j)), "r"(r[j].w), "r"(__float_as_uint(tnv[j])), "r"(k.nthreads * 8u)
                         : "memory");
        t.sp += npush;
    } else
#endif
    if (hitmask) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if ((hitmask >> j) & 1u) {
                const int d = min(t.sp + 1, WIDE_STACK - 1);
                overflow |= (unsigned)(t.sp + 1 >= WIDE_STACK);
                wstk_store(k, d, make_uint2(r[j].w, __float_as_uint(tnv[j])));
                t.sp = d;
            }
        }
    }
#endif
#else
#pragma unroll
    for (int j = 0; j < 8; j++) {
        if (r[j].w == WIDE_EMPTY) break;
        const float tnx = __fmaf_rn(__uint_as_float(__byte_perm(r[j].x, 0x4B000000u, nx)), t.A.x, t.B.x);
        const float tfx = __fmaf_rn(__uint_as_float(__byte_perm(r[j].x, 0x4B000000u, fx)), t.A.x, t.B.x);
        const float tny = __fmaf_rn(__uint_as_float(__byte_perm(r[j].y, 0x4B000000u, ny)), t.A.y, t.B.y);
        const float tfy = __fmaf_rn(__uint_as_float(__byte_perm(r[j].y, 0x4B000000u, fy)), t.A.y, t.B.y);
        const float tnz = __fmaf_rn(__uint_as_float(__byte_perm(r[j].z, 0x4B000000u, nz)), t.A.z, t.B.z);
        const float tfz = __fmaf_rn(__uint_as_float(__byte_perm(r[j].z, 0x4B000000u, fz)), t.A.z, t.B.z);
        const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));
        const float tf = fminf(fminf(tfx, tfy), tfz);
#if WIDE_PRED_PUSH
        // branch-free form: every lane that tests child j runs the same dozen instructions and the
        // store is predicated (r02d profile: the push under a divergent branch was 22 % of all warp
        // instructions at 3 of 32 lanes)
        const bool hitj = tn <= tf && !(tn > t.dist);
        const bool nearer = hitj && tn < best_t;
        const unsigned pr = nearer ? best : r[j].w;
        const float pt = nearer ? best_t : tn;
        best = nearer ? r[j].w : best;
        best_t = nearer ? tn : best_t;
        const bool push = hitj && pr != WIDE_NONE;
        const int dpt = min(t.sp + 1, WIDE_STACK - 1);
        overflow |= (unsigned)(push && t.sp + 1 >= WIDE_STACK);
#if WIDE_STS_ASM
        const unsigned sa = k.sh32 + (unsigned)dpt * (k.nthreads * 8u);
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %0, 0;\n\t@p st.shared.v2.u32 [%1], {%2, %3};\n\t}"
                     :: "r"((unsigned)(push && dpt < WIDE_SSTACK)), "r"(sa), "r"(pr), "r"(__float_as_uint(pt)) : "memory");
        if (push && dpt >= WIDE_SSTACK) k.lo[dpt - WIDE_SSTACK] = make_uint2(pr, __float_as_uint(pt));
#else
        uint2 *slot = wstk_ptr(k, dpt);
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %0, 0;\n\t@p st.v2.u32 [%1], {%2, %3};\n\t}"
                     :: "r"((unsigned)push), "l"(slot), "r"(pr), "r"(__float_as_uint(pt)) : "memory");
#endif
        t.sp = push ? dpt : t.sp;
#else
        if (tn <= tf && !(tn > t.dist)) {
            wide_prefetch_child(w, r[j].w);
            unsigned pr = r[j].w;
            float pt = tn;
            if (tn < best_t) {       // new nearest: the previous nearest goes to the stack
                pr = best; pt = best_t;
                best = r[j].w; best_t = tn;
            }
            if (pr != WIDE_NONE) {
                // a full stack overwrites its top entry and raises the overflow flag (the call fails)
                const int d = min(t.sp + 1, WIDE_STACK - 1);
                overflow |= (unsigned)(t.sp + 1 >= WIDE_STACK);
                wstk_store(k, d, make_uint2(pr, __float_as_uint(pt)));
                t.sp = d;
            }
        }
#endif
    }
#endif  // WIDE_TWO_PASS
#if WIDE_TWO_PASS && !WIDE_EMPTY_CHECK
    // memory safety of the untested empty records, whatever the ray: an empty record's index never becomes the
    // current node (here) nor survives a pop (wtrav_pop / itrav_pop)
    if (best == WIDE_EMPTY) best = WIDE_NONE;
#endif
    if (best != WIDE_NONE) t.cur = best;
    return best == WIDE_NONE;
}

template <bool COUNT>
__device__ __forceinline__ void wtrav_group_step(WTrav &t, const WideView &w, const WStackRef &k,
                                                 unsigned &overflow, uint32_t &ngroup) {
    if (wtrav_group_core<COUNT>(t, w, k, overflow, ngroup)) wtrav_pop(t, k);
}

// Leaf round with the triangle tests SPREAD over the warp (WIDE_LEAF_SPREAD, default on): the
// (ray, triangle) pairs of all lanes that hold a leaf are dealt one per lane, so a round costs one
// memory round trip and one pass through the test at ~2x the lanes instead of up to three passes at
// the 5-8 lanes that hold a leaf (s19 capture: the three test blocks are 6 % of the instructions but
// 24 % of the stall samples, at 5-8 active lanes).  Each owner then folds its results in slot order
// with the serial rule, so the outcome is the serial one bit for bit.
#ifndef WIDE_LEAF_SPREAD
#define WIDE_LEAF_SPREAD 1   // r2b (C3, 1e7 photons): 27.15 -> 26.64 ms, with WIDE_TWO_PASS 26.40 ms
#endif
#define WSPREAD_PAIRS 96          // 32 lanes x leaves of <= 3 triangles; longer leaves take the serial path
struct WSpread {
    unsigned char *own;           // [WSPREAD_PAIRS] lane that owns pair q        (per warp, shared memory)
    float *res;                   // [WSPREAD_PAIRS] distance of pair q, +inf = no hit
};
#define WSPREAD_BYTES_PER_WARP (WSPREAD_PAIRS * 5)

template <bool COUNT>
__device__ __forceinline__ void wtrav_leaf_spread(WTrav &t, bool lw, const WideView &w, const float3 &o,
                                                  const float3 &d, const WStackRef &k, const WSpread &sp,
                                                  uint32_t &ntri) {
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt = (1u << lane) - 1u;
    const unsigned cnt = lw ? ((t.cur >> 28) & 7u) + 1u : 0u;
    if (__any_sync(0xffffffffu, cnt > 3u)) {          // long leaves (CHROMA_WIDE_FAT): serial form
        if (lw) wtrav_leaf_step<COUNT>(t, w, o, d, k, ntri);
        return;
    }
    const unsigned b0 = __ballot_sync(0xffffffffu, cnt & 1u), b1 = __ballot_sync(0xffffffffu, cnt & 2u);
    const unsigned off = __popc(b0 & lt) + 2u * __popc(b1 & lt);
    const unsigned total = __popc(b0) + 2u * __popc(b1);
    const unsigned first = t.cur & 0x0FFFFFFFu;
    for (unsigned j = 0; j < cnt; j++) sp.own[off + j] = (unsigned char)lane;
    __syncwarp();
    for (unsigned base = 0; base < total; base += 32u) {
        const unsigned q = base + lane;
        const bool active = q < total;
        const unsigned src = active ? sp.own[q] : 0u;
        const unsigned o_off = __shfl_sync(0xffffffffu, off, src);
        const unsigned o_first = __shfl_sync(0xffffffffu, first, src);
        const int o_last = __shfl_sync(0xffffffffu, t.last_slot, src);
        RayCtx rc;
        rc.o = f3(__shfl_sync(0xffffffffu, o.x, src), __shfl_sync(0xffffffffu, o.y, src), __shfl_sync(0xffffffffu, o.z, src));
        rc.d = f3(__shfl_sync(0xffffffffu, d.x, src), __shfl_sync(0xffffffffu, d.y, src), __shfl_sync(0xffffffffu, d.z, src));
        if (active) {
            const int slot = (int)(o_first + (q - o_off));
            float tt = __int_as_float(0x7f800000);
            if (slot != o_last) {
                const float4 *p = w.tris + (size_t)slot * 3;
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
            if (COUNT && slot != t.last_slot) ntri++;
            const float tt = sp.res[off + j];
            if (tt < __int_as_float(0x7f800000)) {
                if (tt < t.dist || (tt == t.dist && t.hit >= 0 && __ldg(w.rank + slot) < __ldg(w.rank + t.hit))) {
                    t.hit = slot;
                    t.dist = tt;
                }
            }
        }
        wtrav_pop(t, k);
    }
    __syncwarp();
}

// Warp-level search round: the lanes of a warp that are searching (`searching`) vote between a
// leaf round (one triangle test each) and a group round (eight box tests each).  Lanes that
// reached a leaf wait until `leaf_thresh` of them have one, or no lane has a group left: r01q
// profile of the unscheduled form showed the triangle test running with 3 of 32 lanes.
template <bool COUNT>
__device__ __forceinline__ void wtrav_round(WTrav &t, bool searching, const WideView &w, const float3 &o,
                                            const float3 &d, const WStackRef &k, int leaf_thresh,
                                            unsigned &overflow, uint32_t &ngroup, uint32_t &ntri,
                                            const WSpread &sp) {
    const bool lw = searching && wtrav_leaf_work(t);
    const bool gw = searching && t.cur != WIDE_NONE && !lw;
    const unsigned mL = __ballot_sync(0xffffffffu, lw);
    const unsigned mG = __ballot_sync(0xffffffffu, gw);
    if (mG == 0u || __popc(mL) >= leaf_thresh) {
#if WIDE_LEAF_SPREAD
        wtrav_leaf_spread<COUNT>(t, lw, w, o, d, k, sp, ntri);
#else
        if (lw) wtrav_leaf_step<COUNT>(t, w, o, d, k, ntri);
#endif
    } else {
        if (gw) wtrav_group_step<COUNT>(t, w, k, overflow, ngroup);
    }
}
