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
    if (i < g.top_nodes) return s_nodes[i];
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
