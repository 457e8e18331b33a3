// traverse.cuh — BVH traversal and primitive tests (device code).
//
// Behavioural contract: chroma/cuda/mesh.h:41-118 (intersect_mesh), mesh.h:15-34
// (intersect_node), intersect.h:25-95 (intersect_triangle), intersect.h:106-147
// (intersect_box), geometry.h:30-47 (get_node).  Results (first-hit triangle and the bits
// of its distance) must equal the reference kernels built with --use_fast_math, so every
// floating-point operation below is written with explicit rounding intrinsics in the
// order the reference's SASS performs them (decoded from the reference cubin for sm_100a):
//   cross(a,b).x = fma(a.y, b.z, -(a.z*b.y))       (first product fused, second rounded)
//   dot(a,b)     = fma(a.z, b.z, fma(a.x, b.x, a.y*b.y))
//   f = (float)(1.0/(double)a)  == __frcp_rn(a)    (double rounding is innocuous: 53 >= 2*24+2)
//   (double)x < -1e-6  <=> x < -0x1.0c6f7ap-20f ... (float thresholds below)
// Data layout differs from the reference (leaf slots, packed triangles, shared-memory top
// of the tree); the visit order and every accept/prune decision do not.
#pragma once
#include "common.cuh"

#define CHROMA_STACK_WORDS 192

// thresholds equivalent to the reference's double-precision comparisons on float values
//   (double)x < -1e-6      <=>  x < CHROMA_F_NEG_EPS   (smallest float >= -1e-6)
//   (double)x > 1.0+1e-6   <=>  x > CHROMA_F_ONE_EPS   (largest float  <= 1.000001)
//   (double)x > 1e-6       <=>  x > CHROMA_F_POS_EPS   (largest float  <= 1e-6)
// (bit patterns asserted against numpy in tests/test_host_logic.py)
#define CHROMA_F_NEG_EPS __int_as_float(0xb58637bd)
#define CHROMA_F_ONE_EPS __int_as_float(0x3f800008)
#define CHROMA_F_POS_EPS __int_as_float(0x358637bd)

__device__ __forceinline__ float3 f3(float x, float y, float z) { return make_float3(x, y, z); }

// reference contraction patterns ---------------------------------------------------------
__device__ __forceinline__ float ref_dot(const float3 &a, const float3 &b) {
    return __fmaf_rn(a.z, b.z, __fmaf_rn(a.x, b.x, __fmul_rn(a.y, b.y)));
}
__device__ __forceinline__ float3 ref_cross(const float3 &a, const float3 &b) {
    return f3(__fmaf_rn(a.y, b.z, -__fmul_rn(a.z, b.y)),
              __fmaf_rn(a.z, b.x, -__fmul_rn(a.x, b.z)),
              __fmaf_rn(a.x, b.y, -__fmul_rn(a.y, b.x)));
}
__device__ __forceinline__ float3 ref_sub(const float3 &a, const float3 &b) {
    return f3(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z));
}

struct RayCtx {
    float3 o, d;      // origin, unit direction
    float3 nod, inv;  // -o/d and 1/d with the reference's approximate division
    bool fx, fy, fz;  // isfinite(inv.*)
};

__device__ __forceinline__ void ray_setup(RayCtx &r, const float3 &o, const float3 &d) {
    r.o = o;
    r.d = d;
    // mesh.h:52-53: -origin / direction and 1.0f / direction (div.approx.ftz)
    r.inv = f3(__fdividef(1.0f, d.x), __fdividef(1.0f, d.y), __fdividef(1.0f, d.z));
    r.nod = f3(__fdividef(-o.x, d.x), __fdividef(-o.y, d.y), __fdividef(-o.z, d.z));
    r.fx = isfinite(r.inv.x);
    r.fy = isfinite(r.inv.y);
    r.fz = isfinite(r.inv.z);
}

// geometry.h:30-47 get_node + intersect.h:106-147 intersect_box on the packed node
__device__ __forceinline__ bool node_box_hit(const GeomView &g, const RayCtx &r, const uint4 &n,
                                             float &tmin_out) {
    float tmin = 0.0f, tmax = __int_as_float(0x7f800000);
    if (r.fx) {
        float lo = __fmaf_rn((float)(n.x & 0xFFFFu), g.world_scale, g.world_origin.x);
        float hi = __fmaf_rn((float)(n.x >> 16), g.world_scale, g.world_origin.x);
        float t0 = __fmaf_rn(lo, r.inv.x, r.nod.x);
        float t1 = __fmaf_rn(hi, r.inv.x, r.nod.x);
        tmin = fmaxf(tmin, fminf(t0, t1));
        tmax = fminf(tmax, fmaxf(t0, t1));
    }
    if (r.fy) {
        float lo = __fmaf_rn((float)(n.y & 0xFFFFu), g.world_scale, g.world_origin.y);
        float hi = __fmaf_rn((float)(n.y >> 16), g.world_scale, g.world_origin.y);
        float t0 = __fmaf_rn(lo, r.inv.y, r.nod.y);
        float t1 = __fmaf_rn(hi, r.inv.y, r.nod.y);
        tmin = fmaxf(tmin, fminf(t0, t1));
        tmax = fminf(tmax, fmaxf(t0, t1));
    }
    if (r.fz) {
        float lo = __fmaf_rn((float)(n.z & 0xFFFFu), g.world_scale, g.world_origin.z);
        float hi = __fmaf_rn((float)(n.z >> 16), g.world_scale, g.world_origin.z);
        float t0 = __fmaf_rn(lo, r.inv.z, r.nod.z);
        float t1 = __fmaf_rn(hi, r.inv.z, r.nod.z);
        tmin = fmaxf(tmin, fminf(t0, t1));
        tmax = fminf(tmax, fmaxf(t0, t1));
    }
    tmin_out = tmin;
    return !(tmin > tmax);
}

// packed triangle record: 3 x float4 per leaf slot
//   a = (v0.x, v0.y, v0.z, v1.x)  b = (v1.y, v1.z, v2.x, v2.y)
//   c = (v2.z, material_code, triangle_id, solid_id)   (last three are bit patterns)
struct TriRec {
    float3 v0, v1, v2;
    uint32_t code, id, solid;
};

__device__ __forceinline__ void load_tri(const GeomView &g, uint32_t slot, TriRec &t) {
    const float4 *p = g.tris + (size_t)slot * 3;
    float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    t.v0 = f3(a.x, a.y, a.z);
    t.v1 = f3(a.w, b.x, b.y);
    t.v2 = f3(b.z, b.w, c.x);
    t.code = __float_as_uint(c.y);
    t.id = __float_as_uint(c.z);
    t.solid = __float_as_uint(c.w);
}

// global triangle id -> solid id (placement in add order): last tri_base <= tri
__device__ __forceinline__ uint32_t inst_solid_of_triangle(const uint32_t *tri_base, uint32_t ninst, uint32_t tri) {
    uint32_t lo = 0, hi = ninst;   // tri_base[lo] <= tri < tri_base[hi]
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(tri_base + mid) <= tri) lo = mid; else hi = mid;
    }
    return lo;
}

// solid id of a triangle of the flattened mesh (chroma/cuda/daq.cu:53 solid_map[triangle_id]): the flat
// engine keeps it in the triangle record, the two-level geometry finds the placement whose triangle
// range holds the id.  false: no such triangle.
__device__ __forceinline__ bool solid_of_triangle(const float4 *tris, const int32_t *tri_slot,
                                                  const uint32_t *inst_tri_base, uint32_t ninst,
                                                  uint32_t ntriangles, int tri, uint32_t &solid) {
    if (tri < 0 || (uint32_t)tri >= ntriangles) return false;
    if (inst_tri_base) {
        solid = inst_solid_of_triangle(inst_tri_base, ninst, (uint32_t)tri);
        return true;
    }
    const int slot = tri_slot[tri];
    if (slot < 0) return false;
    solid = __float_as_uint(__ldg(tris + (size_t)slot * 3 + 2).w);
    return true;
}

// intersect.h:25-95 (Moller-Trumbore as compiled for the reference)
__device__ __forceinline__ bool triangle_hit(const RayCtx &r, const float3 &v0, const float3 &v1,
                                             const float3 &v2, float &distance) {
    float3 e1 = ref_sub(v1, v0);
    float3 e2 = ref_sub(v2, v0);
    float3 h = ref_cross(r.d, e2);
    float a = ref_dot(e1, h);
    if (a > -1.1920928955078125e-07f && a < 1.1920928955078125e-07f) return false;
    float f = __frcp_rn(a);
    float3 s = ref_sub(r.o, v0);
    float u = __fmul_rn(f, ref_dot(s, h));
    // (double)u < -1e-6 || (double)u > 1.0+1e-6
    if (u < CHROMA_F_NEG_EPS || u > CHROMA_F_ONE_EPS) return false;
    float3 q = ref_cross(s, e1);
    float v = __fmul_rn(f, ref_dot(r.d, q));
    if (v < CHROMA_F_NEG_EPS || __fadd_rn(u, v) > CHROMA_F_ONE_EPS) return false;
    float t = __fmul_rn(f, ref_dot(e2, q));
    if (t > CHROMA_F_POS_EPS && t < __int_as_float(0x7f800000)) {
        distance = t;
        return true;
    }
    return false;
}

__device__ __forceinline__ uint4 fetch_node(const GeomView &g, const uint4 *s_nodes, uint32_t i) {
    if (s_nodes != nullptr && i < g.top_nodes) return s_nodes[i];  // NULL: no shared-memory copy
    return __ldg(g.nodes + i);
}

// mesh.h:41-118.  Returns the hit leaf slot (-1 none); `best` receives the distance
// (-1.0f when nothing is hit, as the reference leaves min_distance).  `last_slot` is the
// slot of the triangle to be skipped (mesh.h:82), -1 for none.  COUNT enables the visit
// counters of mesh.h:65-66.  `overflow` is set when the stack would exceed its size.
template <bool COUNT>
__device__ __forceinline__ int intersect_mesh_slot(const GeomView &g, const uint4 *s_nodes,
                                                   const RayCtx &r, int last_slot, float &best,
                                                   unsigned *stack, uint32_t *ncount,
                                                   uint32_t *tcount, bool &overflow) {
    int hit = -1;
    float dist = __int_as_float(0x7f800000);  // +inf until a triangle is accepted
    uint32_t nc = 0, tc = 0;
    best = -1.0f;

    uint4 root = fetch_node(g, s_nodes, 0);
    if (COUNT) nc++;
    float tmin;
    if (!node_box_hit(g, r, root, tmin)) {
        if (COUNT) { *ncount = nc; *tcount = tc; }
        return -1;
    }
    int sp = 0;
    stack[0] = root.w;
    while (sp >= 0) {
        unsigned w = stack[sp--];
        unsigned first = w & 0x0FFFFFFFu;
        unsigned end = first + (w >> 28);
        if (first >= end) continue;
        uint4 nxt = fetch_node(g, s_nodes, first);
        for (unsigned i = first; i < end; i++) {
            uint4 n = nxt;
            if (i + 1 < end) nxt = fetch_node(g, s_nodes, i + 1);
            if (COUNT) nc++;
            if (node_box_hit(g, r, n, tmin) && !(tmin > dist)) {
                if ((n.w >> 28) == 0) {
                    int slot = (int)n.w;
                    if (slot != last_slot) {
                        if (COUNT) tc++;
                        const float4 *p = g.tris + (size_t)slot * 3;
                        float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
                        float t;
                        if (triangle_hit(r, f3(a.x, a.y, a.z), f3(a.w, b.x, b.y),
                                         f3(b.z, b.w, c.x), t)) {
                            if (t < dist) {  // first accepted hit always passes (dist=+inf)
                                hit = slot;
                                dist = t;
                            }
                        }
                    }
                } else {
                    if (sp + 1 >= CHROMA_STACK_WORDS) {
                        overflow = true;
                    } else {
                        stack[++sp] = n.w;
                    }
                }
            }
        }
    }
    if (hit >= 0) best = dist;
    if (COUNT) { *ncount = nc; *tcount = tc; }
    return hit;
}

// ---------------------------------------------------------------------------------------
// Warp-scheduled form of the same traversal (v2 kernels).
//
// intersect_mesh_slot() above runs one ray per lane to completion: lanes of a warp diverge
// on the sibling-group length, on the leaf test nested in the group loop and on the total
// trip count (r01 profile: 7 of 32 lanes active in the node test).  Here the traversal is
// cut into two uniform micro-steps that a warp executes phase by phase for all lanes that
// are ready for it:
//   trav_node_step  — test ONE node (pop a sibling group first if the current one is used up);
//                     a leaf whose box passes is parked in `leaf` instead of being tested
//   trav_leaf_step  — run the triangle test of the parked leaf
// A lane's own sequence of node tests, leaf tests and prune decisions is exactly that of
// mesh.h:41-118 (a lane with a parked leaf does no node test until the leaf is resolved, so
// `dist` is always up to date when the next sibling is pruned against it): the results
// stay bit-identical, only the interleaving between lanes changes.
// ---------------------------------------------------------------------------------------
struct Trav {
    float3 inv, nod;   // 1/d and -o/d (approximate division, as the reference)
    unsigned fin;      // bit k set: inv.k is finite; bit 3: ray needs the parallel-axis guard
    unsigned cur, end; // rest of the sibling group being scanned
    int sp;            // top of the pending-group stack (-1 empty)
    float dist;        // best distance so far (+inf none)
    int hit;           // best leaf slot so far (-1 none)
    int last_slot;     // slot excluded from the search (mesh.h:82)
    int leaf;          // parked leaf slot (-1 none)
};

__device__ __forceinline__ bool trav_node_work(const Trav &t) {
    return t.leaf < 0 && (t.cur < t.end || t.sp >= 0);
}
__device__ __forceinline__ bool trav_done(const Trav &t) {
    return t.leaf < 0 && t.cur >= t.end && t.sp < 0;
}

__device__ __forceinline__ bool trav_box(const GeomView &g, const Trav &t, const uint4 &n, float &tmin_out) {
    float tmin = 0.0f, tmax = __int_as_float(0x7f800000);
    if (t.fin & 1u) {
        float lo = __fmaf_rn((float)(n.x & 0xFFFFu), g.world_scale, g.world_origin.x);
        float hi = __fmaf_rn((float)(n.x >> 16), g.world_scale, g.world_origin.x);
        float t0 = __fmaf_rn(lo, t.inv.x, t.nod.x), t1 = __fmaf_rn(hi, t.inv.x, t.nod.x);
        tmin = fmaxf(tmin, fminf(t0, t1));
        tmax = fminf(tmax, fmaxf(t0, t1));
    }
    if (t.fin & 2u) {
        float lo = __fmaf_rn((float)(n.y & 0xFFFFu), g.world_scale, g.world_origin.y);
        float hi = __fmaf_rn((float)(n.y >> 16), g.world_scale, g.world_origin.y);
        float t0 = __fmaf_rn(lo, t.inv.y, t.nod.y), t1 = __fmaf_rn(hi, t.inv.y, t.nod.y);
        tmin = fmaxf(tmin, fminf(t0, t1));
        tmax = fminf(tmax, fmaxf(t0, t1));
    }
    if (t.fin & 4u) {
        float lo = __fmaf_rn((float)(n.z & 0xFFFFu), g.world_scale, g.world_origin.z);
        float hi = __fmaf_rn((float)(n.z >> 16), g.world_scale, g.world_origin.z);
        float t0 = __fmaf_rn(lo, t.inv.z, t.nod.z), t1 = __fmaf_rn(hi, t.inv.z, t.nod.z);
        tmin = fmaxf(tmin, fminf(t0, t1));
        tmax = fminf(tmax, fmaxf(t0, t1));
    }
    tmin_out = tmin;
    return !(tmin > tmax);
}

// Parallel-axis guard.  The reference skips the slab test of an axis whose 1/d is not finite
// (intersect.h:115-139) and its -o/d + x/d form loses all precision when |d.k| is tiny: such a
// ray passes the box test of every node regardless of that axis and walks a large part of the
// whole tree (seconds for one ray in the 170 M-triangle detector; uniform_sphere() returns
// exactly (0,0,+-1) with probability ~3e-8 per call).  None of the extra boxes can contain a
// triangle the ray hits, so rejecting them changes no result: along a guarded axis the ray
// stays within |d.k| * (path bound) of o.k, and boxes farther than that (plus one quantum)
// are dropped before the reference test is applied.
#define CHROMA_GUARD_EPS 1e-4f
__device__ __forceinline__ bool trav_guard_reject(const GeomView &g, const uint4 &n, const float3 &o,
                                                  const float3 &d) {
    // path bound: L1 distance from o to the farthest corner of the world box
    const float ext = 65535.0f * g.world_scale;
    float D = fmaxf(fabsf(o.x - g.world_origin.x), fabsf(o.x - (g.world_origin.x + ext))) +
              fmaxf(fabsf(o.y - g.world_origin.y), fabsf(o.y - (g.world_origin.y + ext))) +
              fmaxf(fabsf(o.z - g.world_origin.z), fabsf(o.z - (g.world_origin.z + ext)));
    const float pad = 2.0f * g.world_scale;
    if (fabsf(d.x) < CHROMA_GUARD_EPS) {
        float r = fabsf(d.x) * D + pad;
        float lo = (float)(n.x & 0xFFFFu) * g.world_scale + g.world_origin.x;
        float hi = (float)(n.x >> 16) * g.world_scale + g.world_origin.x;
        if (lo > o.x + r || hi < o.x - r) return true;
    }
    if (fabsf(d.y) < CHROMA_GUARD_EPS) {
        float r = fabsf(d.y) * D + pad;
        float lo = (float)(n.y & 0xFFFFu) * g.world_scale + g.world_origin.y;
        float hi = (float)(n.y >> 16) * g.world_scale + g.world_origin.y;
        if (lo > o.y + r || hi < o.y - r) return true;
    }
    if (fabsf(d.z) < CHROMA_GUARD_EPS) {
        float r = fabsf(d.z) * D + pad;
        float lo = (float)(n.z & 0xFFFFu) * g.world_scale + g.world_origin.z;
        float hi = (float)(n.z >> 16) * g.world_scale + g.world_origin.z;
        if (lo > o.z + r || hi < o.z - r) return true;
    }
    return false;
}

// mesh.h:47-63: root test and initial stack
__device__ __forceinline__ void trav_begin(Trav &t, const GeomView &g, const uint4 *s_nodes,
                                           const float3 &o, const float3 &d, int last_slot) {
    t.inv = f3(__fdividef(1.0f, d.x), __fdividef(1.0f, d.y), __fdividef(1.0f, d.z));
    t.nod = f3(__fdividef(-o.x, d.x), __fdividef(-o.y, d.y), __fdividef(-o.z, d.z));
    t.fin = (isfinite(t.inv.x) ? 1u : 0u) | (isfinite(t.inv.y) ? 2u : 0u) | (isfinite(t.inv.z) ? 4u : 0u);
    if (fminf(fminf(fabsf(d.x), fabsf(d.y)), fabsf(d.z)) < CHROMA_GUARD_EPS) t.fin |= 8u;
    t.sp = -1;
    t.dist = __int_as_float(0x7f800000);
    t.hit = -1;
    t.last_slot = last_slot;
    t.leaf = -1;
    t.cur = t.end = 0;
    uint4 root = fetch_node(g, s_nodes, 0);
    float tmin;
    if (trav_box(g, t, root, tmin)) {
        t.cur = root.w & 0x0FFFFFFFu;
        t.end = t.cur + (root.w >> 28);
    }
}

// one node test (caller guarantees trav_node_work(t))
template <bool COUNT>
__device__ __forceinline__ void trav_node_step(Trav &t, const GeomView &g, const uint4 *s_nodes,
                                               const float3 &o, const float3 &d, unsigned *stack,
                                               bool &overflow, uint32_t &ncount) {
    if (t.cur >= t.end) {
        unsigned w = stack[t.sp--];
        t.cur = w & 0x0FFFFFFFu;
        t.end = t.cur + (w >> 28);
    }
    uint4 n = fetch_node(g, s_nodes, t.cur);
    t.cur++;
    if (COUNT) ncount++;
    if (!COUNT && (t.fin & 8u) && trav_guard_reject(g, n, o, d)) return;  // counters keep the reference's visits
    float tmin;
    if (trav_box(g, t, n, tmin) && !(tmin > t.dist)) {
        if ((n.w >> 28) == 0) {
            if ((int)n.w != t.last_slot) t.leaf = (int)n.w;
        } else if (t.sp + 1 >= CHROMA_STACK_WORDS) {
            overflow = true;
        } else {
            stack[++t.sp] = n.w;
        }
    }
}

// the parked leaf's triangle test (caller guarantees t.leaf >= 0)
__device__ __forceinline__ void trav_leaf_step(Trav &t, const GeomView &g, const float3 &o,
                                               const float3 &d) {
    const float4 *p = g.tris + (size_t)t.leaf * 3;
    float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    RayCtx r;
    r.o = o;
    r.d = d;
    float tt;
    if (triangle_hit(r, f3(a.x, a.y, a.z), f3(a.w, b.x, b.y), f3(b.z, b.w, c.x), tt)) {
        if (tt < t.dist) {
            t.hit = t.leaf;
            t.dist = tt;
        }
    }
    t.leaf = -1;
}

// ---------------------------------------------------------------------------------------
// Latency-oriented variant of the micro-steps (trace_kernel).  Same tests in the same order;
// what changes is when the loads are issued:
//   * the node that will be tested next is already in registers (`na`): it is requested as
//     soon as its index is known — right after a pop, or one sibling ahead — so its latency
//     overlaps the arithmetic of the current test, the triangle test and the warp scheduling;
//   * when an inner node is pushed, the first line(s) of its child group are prefetched into
//     L2, so the pop that comes later finds them on chip instead of in HBM;
//   * the box test has a branch-free fast path for rays with three finite 1/d components and
//     converts the 16-bit bounds with PRMT + FADD (exact) instead of I2F.
// ---------------------------------------------------------------------------------------
#ifndef CHROMA_PREFETCH_MODE
#define CHROMA_PREFETCH_MODE 2   // 0 none, 1 into L1, 2 into L2
#endif
__device__ __forceinline__ void prefetch_l2(const void *p) {
#if CHROMA_PREFETCH_MODE == 2
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#elif CHROMA_PREFETCH_MODE == 1
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#endif
}

// exact (float)(16-bit field) without I2F: 0x4B000000 | v is 8388608 + v as a float
__device__ __forceinline__ float u16lo_f(uint32_t x) {
    return __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7610)) - 8388608.0f;
}
__device__ __forceinline__ float u16hi_f(uint32_t x) {
    return __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7632)) - 8388608.0f;
}

__device__ __forceinline__ bool trav_box_fast(const GeomView &g, const Trav &t, const uint4 &n, float &tmin_out) {
    // all three axes finite: intersect.h:115-139 without the per-axis branches
    float lo = __fmaf_rn(u16lo_f(n.x), g.world_scale, g.world_origin.x);
    float hi = __fmaf_rn(u16hi_f(n.x), g.world_scale, g.world_origin.x);
    float t0 = __fmaf_rn(lo, t.inv.x, t.nod.x), t1 = __fmaf_rn(hi, t.inv.x, t.nod.x);
    float tmin = fmaxf(0.0f, fminf(t0, t1));
    float tmax = fminf(__int_as_float(0x7f800000), fmaxf(t0, t1));
    lo = __fmaf_rn(u16lo_f(n.y), g.world_scale, g.world_origin.y);
    hi = __fmaf_rn(u16hi_f(n.y), g.world_scale, g.world_origin.y);
    t0 = __fmaf_rn(lo, t.inv.y, t.nod.y); t1 = __fmaf_rn(hi, t.inv.y, t.nod.y);
    tmin = fmaxf(tmin, fminf(t0, t1));
    tmax = fminf(tmax, fmaxf(t0, t1));
    lo = __fmaf_rn(u16lo_f(n.z), g.world_scale, g.world_origin.z);
    hi = __fmaf_rn(u16hi_f(n.z), g.world_scale, g.world_origin.z);
    t0 = __fmaf_rn(lo, t.inv.z, t.nod.z); t1 = __fmaf_rn(hi, t.inv.z, t.nod.z);
    tmin = fmaxf(tmin, fminf(t0, t1));
    tmax = fminf(tmax, fmaxf(t0, t1));
    tmin_out = tmin;
    return !(tmin > tmax);
}

struct TravL : Trav {
    uint4 na;  // node `cur`, already loaded (valid while cur < end)
};

__device__ __forceinline__ void travl_begin(TravL &t, const GeomView &g, const float3 &o,
                                            const float3 &d, int last_slot) {
    trav_begin(t, g, nullptr, o, d, last_slot);
    if (t.cur < t.end) t.na = __ldg(g.nodes + t.cur);
}

// Pending-group stack: the first CHROMA_SSTACK entries of every thread live in shared memory
// (entry-major, so a warp's accesses to one depth hit 32 different banks); deeper entries
// spill to the thread's local array.  r01 profile: with the whole stack in local memory only
// 62 % of the pops hit L1 (1024 threads x 768 B per SM) and the pop was 14 % of all stall
// samples, besides evicting BVH nodes from L1.
#define CHROMA_SSTACK 32
struct StackRef {
    unsigned *sh;   // shared: sh[depth * nthreads + tid]
    unsigned *lo;   // local overflow: lo[depth - CHROMA_SSTACK]
    unsigned nthreads;
};
__device__ __forceinline__ void stk_store(const StackRef &k, int depth, unsigned w) {
    if (depth < CHROMA_SSTACK) k.sh[depth * k.nthreads] = w;
    else k.lo[depth - CHROMA_SSTACK] = w;
}
__device__ __forceinline__ unsigned stk_load(const StackRef &k, int depth) {
    return depth < CHROMA_SSTACK ? k.sh[depth * k.nthreads] : k.lo[depth - CHROMA_SSTACK];
}

// A search micro-step in four parts, ordered so that every load is issued before the first
// instruction that has to wait for an earlier one (the SM issues in order):
//   travl_test   box test of the node in registers, push / park decision, sibling look-ahead
//   travl_fetch_leaf   request the parked leaf's triangle record
//   travl_pop    group used up: pop the next group and request its first node
//   travl_hit    Moller-Trumbore on the requested record (waits for it; the pop's loads fly)
struct LeafRec { float4 a, b, c; };

__device__ __forceinline__ void travl_test(TravL &t, const GeomView &g, const float3 &o,
                                           const float3 &d, const StackRef &k, unsigned &overflow) {
    const uint4 n = t.na;
    t.cur++;
    if (t.cur < t.end) t.na = __ldg(g.nodes + t.cur);  // one sibling ahead
    bool pass;
    float tmin;
    if ((t.fin & 15u) == 7u) {
        pass = trav_box_fast(g, t, n, tmin);
    } else {
        pass = !((t.fin & 8u) && trav_guard_reject(g, n, o, d)) && trav_box(g, t, n, tmin);
    }
    if (pass && !(tmin > t.dist)) {
        if ((n.w >> 28) == 0) {
            if ((int)n.w != t.last_slot) t.leaf = (int)n.w;
        } else if (t.sp + 1 >= CHROMA_STACK_WORDS) {
            overflow = 1u;
        } else {
            stk_store(k, ++t.sp, n.w);
            const uint4 *child = g.nodes + (n.w & 0x0FFFFFFFu);
            prefetch_l2(child);
            if ((n.w >> 28) > 4u) prefetch_l2(child + 8);
        }
    }
}
__device__ __forceinline__ void travl_fetch_leaf(const TravL &t, const GeomView &g, LeafRec &r) {
    const float4 *p = g.tris + (size_t)t.leaf * 3;
    r.a = __ldg(p); r.b = __ldg(p + 1); r.c = __ldg(p + 2);
}
__device__ __forceinline__ void travl_pop(TravL &t, const GeomView &g, const StackRef &k) {
    if (t.cur >= t.end && t.sp >= 0) {
        unsigned w = stk_load(k, t.sp--);
        t.cur = w & 0x0FFFFFFFu;
        t.end = t.cur + (w >> 28);
        t.na = __ldg(g.nodes + t.cur);
    }
}
__device__ __forceinline__ void travl_hit(TravL &t, const LeafRec &r, const float3 &o, const float3 &d) {
    RayCtx rc;
    rc.o = o;
    rc.d = d;
    float tt;
    if (triangle_hit(rc, f3(r.a.x, r.a.y, r.a.z), f3(r.a.w, r.b.x, r.b.y), f3(r.b.z, r.b.w, r.c.x), tt)) {
        if (tt < t.dist) {
            t.hit = t.leaf;
            t.dist = tt;
        }
    }
    t.leaf = -1;
}
// the whole micro-step for one lane (caller guarantees travl_node_work(t))
__device__ __forceinline__ void travl_step(TravL &t, const GeomView &g, const float3 &o, const float3 &d,
                                           const StackRef &k, unsigned &overflow) {
    travl_test(t, g, o, d, k, overflow);
    LeafRec r;
    const bool leaf = t.leaf >= 0;
    if (leaf) travl_fetch_leaf(t, g, r);
    travl_pop(t, g, k);
    if (leaf) travl_hit(t, r, o, d);
}

// with the early pop, "has node work" is simply cur < end
__device__ __forceinline__ bool travl_node_work(const TravL &t) { return t.leaf < 0 && t.cur < t.end; }
__device__ __forceinline__ bool travl_done(const TravL &t) { return t.leaf < 0 && t.cur >= t.end; }
