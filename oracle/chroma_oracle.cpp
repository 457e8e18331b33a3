// chroma_oracle.cpp — CPU restatement of the reference hot path (see chroma_oracle.h).
// TEST INFRASTRUCTURE ONLY: never linked into or called from chroma_b200/.
//
// Written against the reference sources under chroma/cuda/ (cited per function).  Plain
// C++ float arithmetic, compiled with -ffp-contract=off; the fused operations the
// reference's own build performs in the triangle test (decoded from its sm_100a SASS) are
// spelled out with fmaf so that index/distance agree with the GPU kernels to the last bit
// wherever the GPU does not use an approximate instruction.
#include "chroma_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <complex>
#include <numeric>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

struct f3 { float x, y, z; };
inline f3 mk(float x, float y, float z) { return f3{x, y, z}; }
inline f3 operator+(f3 a, f3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
inline f3 operator-(f3 a, f3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
inline f3 operator-(f3 a) { return mk(-a.x, -a.y, -a.z); }
inline f3 operator*(f3 a, float c) { return mk(a.x * c, a.y * c, a.z * c); }
inline f3 operator*(float c, f3 a) { return mk(a.x * c, a.y * c, a.z * c); }
inline f3 operator/(f3 a, float c) { return mk(a.x / c, a.y / c, a.z / c); }
// linalg.h:152-174
inline float dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline f3 cross(f3 a, f3 b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline float norm(f3 a) { return sqrtf(dot(a, a)); }
inline f3 normalize(f3 a) { return a / norm(a); }

// the contraction the reference build applies inside intersect_triangle / norm()
inline float dot_fused(f3 a, f3 b) { return fmaf(a.z, b.z, fmaf(a.x, b.x, a.y * b.y)); }
inline f3 cross_fused(f3 a, f3 b) {
    return mk(fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x)));
}

const float PI_F = 3.141592653589793f;       // physical_constants.h:7
const float SPEED_OF_LIGHT = 299.792458f;    // physical_constants.h:5
const float WEIGHT_LOWER_THRESHOLD = 0.0001f; // photon.h:13
const double CHROMA_EPSILON = 1e-6;          // intersect.h:14

enum { NO_HIT = 1 << 0, BULK_ABSORB = 1 << 1, SURFACE_DETECT = 1 << 2, SURFACE_ABSORB = 1 << 3,
       RAYLEIGH_SCATTER = 1 << 4, REFLECT_DIFFUSE = 1 << 5, REFLECT_SPECULAR = 1 << 6,
       SURFACE_REEMIT = 1 << 7, SURFACE_TRANSMIT = 1 << 8, BULK_REEMIT = 1 << 9 };
const uint32_t NAN_ABORT = 1u << 31;
const uint32_t TERMINAL = NO_HIT | BULK_ABSORB | SURFACE_DETECT | SURFACE_ABSORB | NAN_ABORT;
enum { BREAK, CONTINUE, PASS };
enum { SURFACE_DEFAULT, SURFACE_COMPLEX, SURFACE_WLS, SURFACE_DICHROIC };

// ---------------------------------------------------------------------------------------
// RNG: Philox4x32-10 (Salmon et al., SC'11), counter (id, block, epoch, stream), key seed.
// Stands in for curand_uniform on XORWOW (random.h, photon.h call sites).
// ---------------------------------------------------------------------------------------
struct Rng {
    uint32_t k0, k1, c0, c2, c3, blk, r[4], left;
};
inline void philox_block(const Rng &g, uint32_t out[4]) {
    uint32_t c0 = g.c0, c1 = g.blk, c2 = g.c2, c3 = g.c3, k0 = g.k0, k1 = g.k1;
    for (int i = 0; i < 10; i++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
inline void rng_init(Rng &g, uint64_t seed, uint32_t id, uint32_t epoch, uint32_t stream) {
    g.k0 = (uint32_t)seed; g.k1 = (uint32_t)(seed >> 32);
    g.c0 = id; g.c2 = epoch; g.c3 = stream; g.blk = 0; g.left = 0;
}
inline uint32_t rng_u32(Rng &g) {
    if (g.left == 0) { philox_block(g, g.r); g.blk++; g.left = 4; }
    uint32_t x = g.r[4 - g.left];
    g.left--;
    return x;
}
// curand_uniform: (0,1]
inline float curand_uniform(Rng *g) {
    return fmaf((float)rng_u32(*g), 2.3283064365386963e-10f, 1.1641532182693481e-10f);
}

// random.h:9-13
inline float uniform(Rng *s, float low, float high) { return low + curand_uniform(s) * (high - low); }
// random.h:15-23
inline f3 uniform_sphere(Rng *s) {
    float theta = uniform(s, 0.0f, 2 * PI_F);
    float u = uniform(s, -1.0f, 1.0f);
    float c = sqrtf(1.0f - u * u);
    return mk(c * cosf(theta), c * sinf(theta), u);
}
// interpolate.h:32-58
inline float interp(float x, int n, const float *xp, const float *fp) {
    int lower = 0, upper = n - 1;
    if (x <= xp[lower]) return fp[lower];
    if (x >= xp[upper]) return fp[upper];
    while (lower < upper - 1) {
        int half = (lower + upper) / 2;
        if (x < xp[half]) upper = half; else lower = half;
    }
    float df = fp[upper] - fp[lower];
    float dx = xp[upper] - xp[lower];
    return fp[lower] + df * (x - xp[lower]) / dx;
}
// interpolate.h:4-29
inline float interp_idx(float x, int n, const float *xp) {
    int lower = 0, upper = n - 1;
    if (x <= xp[lower]) return lower;
    if (x >= xp[upper]) return upper;
    while (lower < upper - 1) {
        int half = (lower + upper) / 2;
        if (x < xp[half]) upper = half; else lower = half;
    }
    float dx = xp[upper] - xp[lower];
    return lower + 1.0 * (x - xp[lower]) / dx;
}
// random.h:27-31
inline float sample_cdf(Rng *rng, int ncdf, const float *cdf_x, const float *cdf_y) {
    return interp(curand_uniform(rng), ncdf, cdf_y, cdf_x);
}
// random.h:33-55
inline float sample_cdf(Rng *rng, int ncdf, float x0, float delta, const float *cdf_y) {
    float u = curand_uniform(rng);
    int lower = 0, upper = ncdf - 1;
    while (lower < upper - 1) {
        int half = (lower + upper) / 2;
        if (u < cdf_y[half]) upper = half; else lower = half;
    }
    float delta_cdf_y = cdf_y[upper] - cdf_y[lower];
    return x0 + delta * lower + delta * (u - cdf_y[lower]) / delta_cdf_y;
}

// ---------------------------------------------------------------------------------------
// geometry access (geometry.h, geometry_types.h)
// ---------------------------------------------------------------------------------------
const unsigned CHILD_BITS = 28;
const unsigned NCHILD_MASK = 0xFFFFu << CHILD_BITS;
struct Node { f3 lower, upper; unsigned child, nchild; };
struct Triangle { f3 v0, v1, v2; };

struct Geo {
    const oracle_geometry *d;
    std::vector<uint32_t> comp_first;                 // first component of each material
    std::vector<uint64_t> dich_first;                 // first angle row of each surface
    explicit Geo(const oracle_geometry *g) : d(g) {
        comp_first.resize(g->nmaterials ? g->nmaterials : 1, 0);
        uint32_t run = 0;
        for (uint32_t m = 0; m < g->nmaterials; m++) {
            comp_first[m] = run;
            run += g->material_num_comp ? g->material_num_comp[m] : 0;
        }
        dich_first.resize(g->nsurfaces ? g->nsurfaces : 1, 0);
        uint64_t acc = 0;
        for (uint32_t s = 0; s < g->nsurfaces; s++) {
            dich_first[s] = acc;
            acc += g->dichroic_nangles ? g->dichroic_nangles[s] : 0;
        }
    }
    const float *mat(int m, int k) const { return d->material_props + ((size_t)m * 3 + k) * d->nwavelengths; }
    const float *surf(int s, int k) const { return d->surface_props + ((size_t)s * 8 + k) * d->nwavelengths; }
};

// geometry.h:30-47 get_node
inline Node get_node(const oracle_geometry *g, unsigned i) {
    const uint32_t *n = g->nodes + 4 * (size_t)i;
    Node s;
    // world_origin + to_float3(int) * world_scale, contracted to an FMA by the reference build
    s.lower = mk(fmaf((float)(n[0] & 0xFFFF), g->world_scale, g->world_origin[0]),
                 fmaf((float)(n[1] & 0xFFFF), g->world_scale, g->world_origin[1]),
                 fmaf((float)(n[2] & 0xFFFF), g->world_scale, g->world_origin[2]));
    s.upper = mk(fmaf((float)(n[0] >> 16), g->world_scale, g->world_origin[0]),
                 fmaf((float)(n[1] >> 16), g->world_scale, g->world_origin[1]),
                 fmaf((float)(n[2] >> 16), g->world_scale, g->world_origin[2]));
    s.child = n[3] & ~NCHILD_MASK;
    s.nchild = n[3] >> CHILD_BITS;
    return s;
}
// geometry.h:49-60 get_triangle
inline Triangle get_triangle(const oracle_geometry *g, unsigned i) {
    const uint32_t *t = g->triangles + 3 * (size_t)i;
    const float *a = g->vertices + 3 * (size_t)t[0], *b = g->vertices + 3 * (size_t)t[1],
                *c = g->vertices + 3 * (size_t)t[2];
    return Triangle{mk(a[0], a[1], a[2]), mk(b[0], b[1], b[2]), mk(c[0], c[1], c[2])};
}
// geometry.h:62-75 interp_property
inline float interp_property(const oracle_geometry *m, float x, const float *fp) {
    if (x < m->wavelength_start) return fp[0];
    if (x > (m->wavelength_start + (m->nwavelengths - 1) * m->wavelength_step)) return fp[m->nwavelengths - 1];
    int jl = (x - m->wavelength_start) / m->wavelength_step;
    float next = (jl + 1 < (int)m->nwavelengths) ? fp[jl + 1] : fp[jl];  // reference reads fp[n] with weight 0
    return fp[jl] + (x - (m->wavelength_start + jl * m->wavelength_step)) * (next - fp[jl]) / m->wavelength_step;
}

// intersect.h:25-95 intersect_triangle
inline bool intersect_triangle(const f3 &origin, const f3 &direction, const Triangle &tr, float &distance) {
    f3 edge1 = tr.v1 - tr.v0;
    f3 edge2 = tr.v2 - tr.v0;
    f3 h = cross_fused(direction, edge2);
    float a = dot_fused(edge1, h);
    if (a > -1.1920928955078125e-07f && a < 1.1920928955078125e-07f) return false;  // FLT_EPSILON
    float f = (float)(1.0 / (double)a);
    f3 s = origin - tr.v0;
    float u = f * dot_fused(s, h);
    if ((double)u < -CHROMA_EPSILON || (double)u > 1.0 + CHROMA_EPSILON) return false;
    f3 q = cross_fused(s, edge1);
    float v = f * dot_fused(direction, q);
    float upv = u + v;
    if ((double)v < -CHROMA_EPSILON || (double)upv > 1.0 + CHROMA_EPSILON) return false;
    float t = f * dot_fused(edge2, q);
    if ((double)t > CHROMA_EPSILON && t < INFINITY) {
        distance = t;
        return true;
    }
    return false;
}
// intersect.h:106-147 intersect_box
inline bool intersect_box(const f3 &nod, const f3 &inv, const f3 &lo, const f3 &hi, float &distance_to_box) {
    float tmin = 0.0f, tmax = INFINITY, t0, t1;
    if (std::isfinite(inv.x)) {
        t0 = fmaf(lo.x, inv.x, nod.x); t1 = fmaf(hi.x, inv.x, nod.x);
        tmin = fmaxf(tmin, fminf(t0, t1)); tmax = fminf(tmax, fmaxf(t0, t1));
    }
    if (std::isfinite(inv.y)) {
        t0 = fmaf(lo.y, inv.y, nod.y); t1 = fmaf(hi.y, inv.y, nod.y);
        tmin = fmaxf(tmin, fminf(t0, t1)); tmax = fminf(tmax, fmaxf(t0, t1));
    }
    if (std::isfinite(inv.z)) {
        t0 = fmaf(lo.z, inv.z, nod.z); t1 = fmaf(hi.z, inv.z, nod.z);
        tmin = fmaxf(tmin, fminf(t0, t1)); tmax = fminf(tmax, fmaxf(t0, t1));
    }
    if (tmin > tmax) return false;
    distance_to_box = tmin;
    return true;
}
// mesh.h:15-34 intersect_node
inline bool intersect_node(const f3 &nod, const f3 &inv, const Node &node, float min_distance) {
    float distance_to_box;
    if (intersect_box(nod, inv, node.lower, node.upper, distance_to_box)) {
        if (min_distance < 0.0f) return true;
        if (distance_to_box > min_distance) return false;
        return true;
    }
    return false;
}

const int STACK_SIZE = 1000;  // mesh.h:9

// mesh.h:41-118 intersect_mesh (with the count / tri_count counters of :65-66 live)
int intersect_mesh(const f3 &origin, const f3 &direction, const oracle_geometry *g, float &min_distance,
                   int last_hit_triangle, unsigned &count, unsigned &tri_count) {
    int triangle_index = -1;
    float distance;
    min_distance = -1.0f;
    Node root = get_node(g, 0);
    count++;
    // -origin / direction and 1.0f / direction; under --use_fast_math x/y is x * rcp(y)
    f3 inv = mk(1.0f / direction.x, 1.0f / direction.y, 1.0f / direction.z);
    f3 nod = mk(-origin.x * inv.x, -origin.y * inv.y, -origin.z * inv.z);
    if (!intersect_node(nod, inv, root, min_distance)) return -1;

    unsigned child_ptr_stack[STACK_SIZE], nchild_ptr_stack[STACK_SIZE];
    child_ptr_stack[0] = root.child;
    nchild_ptr_stack[0] = root.nchild;
    int curr = 0;
    while (curr >= 0) {
        unsigned first_child = child_ptr_stack[curr], nchild = nchild_ptr_stack[curr];
        curr--;
        for (unsigned i = first_child; i < first_child + nchild; i++) {
            Node node = get_node(g, i);
            count++;
            if (intersect_node(nod, inv, node, min_distance)) {
                if (node.nchild == 0) {  // leaf: wraps one triangle
                    if ((int)node.child != last_hit_triangle) {
                        tri_count++;
                        Triangle t = get_triangle(g, node.child);
                        if (intersect_triangle(origin, direction, t, distance)) {
                            if (triangle_index == -1 || distance < min_distance) {
                                triangle_index = node.child;
                                min_distance = distance;
                            }
                        }
                    }
                } else {
                    curr++;
                    child_ptr_stack[curr] = node.child;
                    nchild_ptr_stack[curr] = node.nchild;
                }
            }
            if (curr >= STACK_SIZE - 1) break;  // "intersect_mesh() aborted; node > tail"
        }
    }
    return triangle_index;
}

}  // namespace

extern "C" int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
extern "C" void oracle_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

// mesh.h:123-151 distance_to_mesh (+ the triangle index)
extern "C" void oracle_intersect(const oracle_geometry *g, uint64_t n, const float *pos, const float *dir,
                                 const int32_t *last_hit, int32_t *out_tri, float *out_dist,
                                 uint32_t *out_counts) {
#pragma omp parallel for schedule(dynamic, 256)
    for (int64_t id = 0; id < (int64_t)n; id++) {
        f3 origin = mk(pos[3 * id], pos[3 * id + 1], pos[3 * id + 2]);
        f3 direction = mk(dir[3 * id], dir[3 * id + 1], dir[3 * id + 2]);
        float nrm = sqrtf(dot_fused(direction, direction));
        float rn = 1.0f / nrm;  // direction /= norm(direction): x * rcp(norm) under fast-math
        direction = mk(direction.x * rn, direction.y * rn, direction.z * rn);
        float distance;
        unsigned count = 0, tri_count = 0;
        int tri = intersect_mesh(origin, direction, g, distance, last_hit ? last_hit[id] : -1, count, tri_count);
        out_tri[id] = tri;
        out_dist[id] = distance;
        if (out_counts) { out_counts[2 * id] = count; out_counts[2 * id + 1] = tri_count; }
    }
}

// Experiment / cross-check: the same tree searched near-first (children of a node are visited
// in order of box entry distance, subtrees farther than the best hit are skipped).  Returns the
// geometric nearest hit; out_counts = (box tests, triangle tests).  Used to size the gain of an
// ordered fast path and to check that it finds the reference's answer (tests/).
extern "C" void oracle_intersect_ordered(const oracle_geometry *g, uint64_t n, const float *pos,
                                         const float *dir, const int32_t *last_hit, int32_t *out_tri,
                                         float *out_dist, uint32_t *out_counts) {
#pragma omp parallel for schedule(dynamic, 256)
    for (int64_t id = 0; id < (int64_t)n; id++) {
        f3 origin = mk(pos[3 * id], pos[3 * id + 1], pos[3 * id + 2]);
        f3 direction = mk(dir[3 * id], dir[3 * id + 1], dir[3 * id + 2]);
        float rn = 1.0f / sqrtf(dot_fused(direction, direction));
        direction = mk(direction.x * rn, direction.y * rn, direction.z * rn);
        f3 inv = mk(1.0f / direction.x, 1.0f / direction.y, 1.0f / direction.z);
        f3 nod = mk(-origin.x * inv.x, -origin.y * inv.y, -origin.z * inv.z);
        int last = last_hit ? last_hit[id] : -1;
        unsigned count = 1, tri_count = 0;
        int best_tri = -1;
        float best = INFINITY;
        struct Ent { float t; unsigned idx; };
        Ent stack[4096];
        int sp = 0;
        Node root = get_node(g, 0);
        float tb;
        if (intersect_box(nod, inv, root.lower, root.upper, tb)) stack[sp++] = Ent{tb, 0u};
        while (sp > 0) {
            Ent e = stack[--sp];
            if (e.t > best) continue;
            Node node = get_node(g, e.idx);
            if (node.nchild == 0) {
                if ((int)node.child == last) continue;
                tri_count++;
                float distance;
                Triangle t = get_triangle(g, node.child);
                if (intersect_triangle(origin, direction, t, distance) && distance < best) {
                    best = distance;
                    best_tri = node.child;
                }
                continue;
            }
            Ent kids[16];
            int nk = 0;
            for (unsigned i = node.child; i < node.child + node.nchild; i++) {
                Node c = get_node(g, i);
                count++;
                if (intersect_box(nod, inv, c.lower, c.upper, tb) && !(tb > best)) kids[nk++] = Ent{tb, i};
            }
            std::sort(kids, kids + nk, [](const Ent &a, const Ent &b) { return a.t > b.t; });  // far first
            for (int k = 0; k < nk && sp < 4096; k++) stack[sp++] = kids[k];
        }
        out_tri[id] = best_tri;
        out_dist[id] = best_tri >= 0 ? best : -1.0f;
        if (out_counts) { out_counts[2 * id] = count; out_counts[2 * id + 1] = tri_count; }
    }
}

extern "C" void oracle_uniform_stream(uint64_t seed, uint32_t id, uint32_t epoch, uint32_t stream,
                                      uint32_t n, float *out) {
    Rng g;
    rng_init(g, seed, id, epoch, stream);
    for (uint32_t i = 0; i < n; i++) out[i] = curand_uniform(&g);
}

// ---------------------------------------------------------------------------------------
// Photon source (include/chroma_b200.h chroma_photons_generate): the host-side bomb of
// chroma/benchmark.py:75-83 / doc/source/example_detector.py:116-121 with the samplers of
// chroma/sample.py:4-57, restated in float on the engine's Philox stream (seed; id, sequence, 3).
// sample.py:17-29: theta ~ U(0,2pi), u ~ U(cos phi,1), c = sqrt(1-u^2), (c cos theta, c sin theta, u);
// sample.py:38-57: rotated onto `direction` (rot, built by the caller as sample.py does);
// benchmark.py:80: pol = normalize(cross(uniform_sphere, dir)); benchmark.py:81: wavelength ~ U(lo,hi).
// ---------------------------------------------------------------------------------------
static inline void source_cone(float u0, float u1, float cos_half, float out[3]) {
    const double a = M_PI * (double)(2.0f * u0);
    const float s = (float)sin(a), c = (float)cos(a);
    const float u = fmaf(u1, 1.0f - cos_half, cos_half);
    const float q = sqrtf(fmaxf(1.0f - u * u, 0.0f));
    out[0] = q * c; out[1] = q * s; out[2] = u;
}

extern "C" void oracle_generate_photons(uint64_t n, uint32_t id0, uint64_t seed, uint32_t sequence,
                                        const float *rot, float cos_half, float wl_lo, float wl_hi,
                                        float *dir, float *pol, float *wavelengths) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)n; i++) {
        Rng g;
        rng_init(g, seed, id0 + (uint32_t)i, sequence, 3u);
        const float u0 = curand_uniform(&g), u1 = curand_uniform(&g), u2 = curand_uniform(&g),
                    u3 = curand_uniform(&g), u4 = curand_uniform(&g);
        float l[3], a[3], d[3], e[3];
        source_cone(u0, u1, cos_half, l);
        for (int k = 0; k < 3; k++)
            d[k] = fmaf(rot[3 * k + 2], l[2], fmaf(rot[3 * k + 1], l[1], rot[3 * k] * l[0]));
        source_cone(u2, u3, -1.0f, a);
        e[0] = fmaf(a[1], d[2], -(a[2] * d[1]));
        e[1] = fmaf(a[2], d[0], -(a[0] * d[2]));
        e[2] = fmaf(a[0], d[1], -(a[1] * d[0]));
        float nn = fmaf(e[2], e[2], fmaf(e[1], e[1], e[0] * e[0]));
        if (!(nn > 1e-12f)) {
            if (fabsf(d[2]) < 0.9f) { e[0] = d[1]; e[1] = -d[0]; e[2] = 0.0f; }
            else { e[0] = 0.0f; e[1] = d[2]; e[2] = -d[1]; }
            nn = fmaf(e[2], e[2], fmaf(e[1], e[1], e[0] * e[0]));
        }
        const float inv = 1.0f / sqrtf(nn);
        for (int k = 0; k < 3; k++) { dir[3 * i + k] = d[k]; pol[3 * i + k] = e[k] * inv; }
        wavelengths[i] = fmaf(u4, wl_hi - wl_lo, wl_lo);
    }
}

// =======================================================================================
// photon physics (photon.h)
// =======================================================================================
namespace {

struct Photon {  // photon.h:15-30
    f3 position, direction, polarization;
    float wavelength, time, weight;
    unsigned history;
    int last_hit_triangle;
};
struct State {   // photon.h:32-47
    bool inside_to_outside;
    f3 surface_normal;
    float refractive_index1, refractive_index2, absorption_length, scattering_length;
    int material1;
    int surface_index;
    float distance_to_boundary;
};

// photon.h:68-75
inline int convert(int c) { return (c & 0x80) ? (int)(0xFFFFFF00u | c) : c; }
// photon.h:77-81
inline float get_theta(const f3 &a, const f3 &b) { return acosf(fmaxf(-1.0f, fminf(1.0f, dot(a, b)))); }
// rotate.h:22-28
inline f3 rotate(const f3 &a, float phi, const f3 &n) {
    float cos_phi = cosf(phi), sin_phi = sinf(phi);
    return a * cos_phi + n * dot(a, n) * (1.0f - cos_phi) + cross(a, n) * sin_phi;
}

// photon.h:83-135
void fill_state(State &s, Photon &p, const Geo &G, unsigned &count, unsigned &tri_count) {
    const oracle_geometry *g = G.d;
    p.last_hit_triangle = intersect_mesh(p.position, p.direction, g, s.distance_to_boundary,
                                         p.last_hit_triangle, count, tri_count);
    if (p.last_hit_triangle == -1) {
        p.history |= NO_HIT;
        return;
    }
    Triangle t = get_triangle(g, p.last_hit_triangle);
    unsigned material_code = g->material_codes[p.last_hit_triangle];
    int inner_material_index = convert(0xFF & (material_code >> 24));
    int outer_material_index = convert(0xFF & (material_code >> 16));
    s.surface_index = convert(0xFF & (material_code >> 8));
    f3 v01 = t.v1 - t.v0;
    f3 v12 = t.v2 - t.v1;
    s.surface_normal = normalize(cross(v01, v12));
    int material1, material2;
    if (dot(s.surface_normal, -p.direction) > 0.0f) {  // outside to inside
        material1 = outer_material_index;
        material2 = inner_material_index;
        s.inside_to_outside = false;
    } else {                                           // inside to outside
        material1 = inner_material_index;
        material2 = outer_material_index;
        s.surface_normal = -s.surface_normal;
        s.inside_to_outside = true;
    }
    s.refractive_index1 = interp_property(g, p.wavelength, G.mat(material1, 0));
    s.refractive_index2 = interp_property(g, p.wavelength, G.mat(material2, 0));
    s.absorption_length = interp_property(g, p.wavelength, G.mat(material1, 1));
    s.scattering_length = interp_property(g, p.wavelength, G.mat(material1, 2));
    s.material1 = material1;
}

// photon.h:137-165
f3 pick_new_direction(f3 axis, float theta, float phi) {
    float cos_theta = cosf(theta), sin_theta = sinf(theta);
    float cos_phi = cosf(phi), sin_phi = sinf(phi);
    float sin_axis_theta = sqrtf(1.0f - axis.z * axis.z);
    float cos_axis_phi, sin_axis_phi;
    if (std::isnan(sin_axis_theta) || sin_axis_theta < 0.00001f) {
        cos_axis_phi = 1.0f;
        sin_axis_phi = 0.0f;
    } else {
        cos_axis_phi = axis.x / sin_axis_theta;
        sin_axis_phi = axis.y / sin_axis_theta;
    }
    float dirx = cos_theta * axis.x + sin_theta * (axis.z * cos_phi * cos_axis_phi - sin_phi * sin_axis_phi);
    float diry = cos_theta * axis.y + sin_theta * (cos_phi * axis.z * sin_axis_phi + sin_phi * cos_axis_phi);
    float dirz = cos_theta * axis.z - sin_theta * cos_phi * sin_axis_theta;
    return mk(dirx, diry, dirz);
}

// photon.h:167-191
void rayleigh_scatter(Photon &p, Rng &rng) {
    float cos_theta = 2.0f * cosf((acosf(1.0f - 2.0f * curand_uniform(&rng)) - 2 * PI_F) / 3.0f);
    if (cos_theta > 1.0f) cos_theta = 1.0f;
    else if (cos_theta < -1.0f) cos_theta = -1.0f;
    float theta = acosf(cos_theta);
    float phi = uniform(&rng, 0.0f, 2.0f * PI_F);
    p.direction = pick_new_direction(p.polarization, theta, phi);
    if (1.0f - fabsf(cos_theta) < 1e-6f) {
        p.polarization = pick_new_direction(p.polarization, PI_F / 2.0f, phi);
    } else {
        p.polarization = p.polarization - cos_theta * p.direction;
    }
    p.direction = p.direction / norm(p.direction);
    p.polarization = p.polarization / norm(p.polarization);
}

// photon.h:193-308
int propagate_to_boundary(Photon &p, State &s, Rng &rng, const Geo &G, bool use_weights, int scatter_first) {
    const oracle_geometry *g = G.d;
    float absorption_distance = -s.absorption_length * logf(curand_uniform(&rng));
    float scattering_distance = -s.scattering_length * logf(curand_uniform(&rng));
    if (use_weights && p.weight > WEIGHT_LOWER_THRESHOLD)
        absorption_distance = 1e30f;
    else
        use_weights = false;

    if (scatter_first == 1) {
        float scatter_prob = 1.0f - expf(-s.distance_to_boundary / s.scattering_length);
        if (scatter_prob > WEIGHT_LOWER_THRESHOLD) {
            int i = 0;
            const int max_i = 1000;
            while (i < max_i && scattering_distance > s.distance_to_boundary) {
                scattering_distance = -s.scattering_length * logf(curand_uniform(&rng));
                i++;
            }
            p.weight *= scatter_prob;
        }
    } else if (scatter_first == -1) {
        float no_scatter_prob = expf(-s.distance_to_boundary / s.scattering_length);
        if (no_scatter_prob > WEIGHT_LOWER_THRESHOLD) {
            int i = 0;
            const int max_i = 1000;
            while (i < max_i && scattering_distance <= s.distance_to_boundary) {
                scattering_distance = -s.scattering_length * logf(curand_uniform(&rng));
                i++;
            }
            p.weight *= no_scatter_prob;
        }
    }

    if (absorption_distance <= scattering_distance) {
        if (absorption_distance <= s.distance_to_boundary) {
            p.time += absorption_distance / (SPEED_OF_LIGHT / s.refractive_index1);
            p.position = p.position + absorption_distance * p.direction;
            unsigned num_comp = g->material_num_comp ? g->material_num_comp[s.material1] : 0;
            if (num_comp == 0) {
                p.last_hit_triangle = -1;
                p.history |= BULK_ABSORB;
                return BREAK;
            }
            const size_t nw = g->nwavelengths;
            unsigned c0 = G.comp_first[s.material1];
            float uniform_sample_comp = curand_uniform(&rng);
            float prob = 0.0f;
            unsigned comp;
            for (comp = 0;; comp++) {
                float comp_abs = interp_property(g, p.wavelength, g->comp_absorption_length + (c0 + comp) * nw);
                prob += s.absorption_length / comp_abs;
                if (uniform_sample_comp < prob || comp + 1 == num_comp) break;
            }
            float uniform_sample_reemit = curand_uniform(&rng);
            float comp_reemit_prob = interp_property(g, p.wavelength, g->comp_reemission_prob + (c0 + comp) * nw);
            if (uniform_sample_reemit < comp_reemit_prob) {
                p.wavelength = sample_cdf(&rng, (int)g->nwavelengths, g->wavelength_start, g->wavelength_step,
                                          g->comp_reemission_wvl_cdf + (c0 + comp) * nw);
                p.time += sample_cdf(&rng, (int)g->ntimes, g->time_start, g->time_step,
                                     g->comp_reemission_time_cdf + (size_t)(c0 + comp) * g->ntimes);
                p.direction = uniform_sphere(&rng);
                p.polarization = cross(uniform_sphere(&rng), p.direction);
                p.polarization = p.polarization / norm(p.polarization);
                p.last_hit_triangle = -1;
                p.history |= BULK_REEMIT;
                return CONTINUE;
            } else {
                p.last_hit_triangle = -1;
                p.history |= BULK_ABSORB;
                return BREAK;
            }
        }
    } else {
        if (scattering_distance <= s.distance_to_boundary) {
            if (use_weights) p.weight *= expf(-scattering_distance / s.absorption_length);
            p.time += scattering_distance / (SPEED_OF_LIGHT / s.refractive_index1);
            p.position = p.position + scattering_distance * p.direction;
            rayleigh_scatter(p, rng);
            p.history |= RAYLEIGH_SCATTER;
            p.last_hit_triangle = -1;
            return CONTINUE;
        }
    }
    if (use_weights) p.weight *= expf(-s.distance_to_boundary / s.absorption_length);
    p.position = p.position + s.distance_to_boundary * p.direction;
    p.time += s.distance_to_boundary / (SPEED_OF_LIGHT / s.refractive_index1);
    return PASS;
}

// photon.h:310-363
void propagate_at_boundary(Photon &p, State &s, Rng &rng) {
    float incident_angle = get_theta(s.surface_normal, -p.direction);
    float refracted_angle = asinf(sinf(incident_angle) * s.refractive_index1 / s.refractive_index2);
    f3 incident_plane_normal = cross(p.direction, s.surface_normal);
    float incident_plane_normal_length = norm(incident_plane_normal);
    if (incident_plane_normal_length < 1e-6f)
        incident_plane_normal = p.polarization;
    else
        incident_plane_normal = incident_plane_normal / incident_plane_normal_length;
    float normal_coefficient = dot(p.polarization, incident_plane_normal);
    float normal_probability = normal_coefficient * normal_coefficient;
    float reflection_coefficient;
    if (curand_uniform(&rng) < normal_probability) {
        reflection_coefficient = -sinf(incident_angle - refracted_angle) / sinf(incident_angle + refracted_angle);
        if ((curand_uniform(&rng) < reflection_coefficient * reflection_coefficient) || std::isnan(refracted_angle)) {
            p.direction = rotate(s.surface_normal, incident_angle, incident_plane_normal);
            p.history |= REFLECT_SPECULAR;
        } else {
            p.direction = rotate(s.surface_normal, PI_F - refracted_angle, incident_plane_normal);
        }
        p.polarization = incident_plane_normal;
    } else {
        reflection_coefficient = tanf(incident_angle - refracted_angle) / tanf(incident_angle + refracted_angle);
        if ((curand_uniform(&rng) < reflection_coefficient * reflection_coefficient) || std::isnan(refracted_angle)) {
            p.direction = rotate(s.surface_normal, incident_angle, incident_plane_normal);
            p.history |= REFLECT_SPECULAR;
        } else {
            p.direction = rotate(s.surface_normal, PI_F - refracted_angle, incident_plane_normal);
        }
        p.polarization = cross(incident_plane_normal, p.direction);
        p.polarization = p.polarization / norm(p.polarization);
    }
}

// photon.h:365-377
int propagate_at_specular_reflector(Photon &p, State &s) {
    float incident_angle = get_theta(s.surface_normal, -p.direction);
    f3 incident_plane_normal = cross(p.direction, s.surface_normal);
    incident_plane_normal = incident_plane_normal / norm(incident_plane_normal);
    p.direction = rotate(s.surface_normal, incident_angle, incident_plane_normal);
    p.history |= REFLECT_SPECULAR;
    return CONTINUE;
}
// photon.h:379-398
int propagate_at_diffuse_reflector(Photon &p, State &s, Rng &rng) {
    float ndotv;
    do {
        p.direction = uniform_sphere(&rng);
        ndotv = dot(p.direction, s.surface_normal);
        if (ndotv < 0.0f) {
            p.direction = -p.direction;
            ndotv = -ndotv;
        }
    } while (!(curand_uniform(&rng) < ndotv));
    p.polarization = cross(uniform_sphere(&rng), p.direction);
    p.polarization = p.polarization / norm(p.polarization);
    p.history |= REFLECT_DIFFUSE;
    return CONTINUE;
}

// complex helpers: cuComplex.h (cuCmulf/cuCdivf/cuCabsf) and cx.h:24-34
typedef std::complex<float> cx;
inline cx cmul(cx a, cx b) { return cx(a.real() * b.real() - a.imag() * b.imag(), a.real() * b.imag() + a.imag() * b.real()); }
inline cx cdiv(cx x, cx y) {  // cuCdivf: scaled to avoid overflow
    float s = fabsf(y.real()) + fabsf(y.imag());
    float oos = 1.0f / s;
    float ars = x.real() * oos, ais = x.imag() * oos, brs = y.real() * oos, bis = y.imag() * oos;
    s = (brs * brs) + (bis * bis);
    oos = 1.0f / s;
    return cx(((ars * brs) + (ais * bis)) * oos, ((ais * brs) - (ars * bis)) * oos);
}
inline float cabs_(cx x) {  // cuCabsf
    float a = fabsf(x.real()), b = fabsf(x.imag()), v, w, t;
    if (a > b) { v = a; w = b; } else { v = b; w = a; }
    t = w / v;
    t = 1.0f + t * t;
    t = v * sqrtf(t);
    if ((v == 0.0f) || (v > 3.402823466e38f) || (w > 3.402823466e38f)) t = v + w;
    return t;
}
inline float carg_(cx x) { return atan2f(x.imag(), x.real()); }
inline cx csqrt_(cx x) {
    float r = sqrtf(cabs_(x));
    float t = carg_(x) / 2.0f;
    return cx(r * cosf(t), r * sinf(t));
}

// one polarisation of photon.h:426-452 (s), :455-482 (p), :484-508 (normal incidence)
void thin_film(cx r12, cx r23, cx t12, cx t23, cx gg, float u, float v, float e, float &R, float &T) {
    float abs_r12 = cabs_(r12), abs_r23 = cabs_(r23), abs_t12 = cabs_(t12), abs_t23 = cabs_(t23);
    float arg_r12 = carg_(r12), arg_r23 = carg_(r23);
    float exp1 = expf(2.0f * v * e);
    float exp2 = 1.0f / exp1;
    float denom = exp1 + abs_r12 * abs_r12 * abs_r23 * abs_r23 * exp2 +
                  2.0f * abs_r12 * abs_r23 * cosf(arg_r23 + arg_r12 + 2.0f * u * e);
    float r = abs_r12 * abs_r12 * exp1 + abs_r23 * abs_r23 * exp2 +
              2.0f * abs_r12 * abs_r23 * cosf(arg_r23 - arg_r12 + 2.0f * u * e);
    R = r / denom;
    float t = gg.real() * abs_t12 * abs_t12 * abs_t23 * abs_t23;
    T = t / denom;
}

// photon.h:400-590
int propagate_complex(Photon &p, State &s, Rng &rng, const Geo &G, int si, bool use_weights) {
    const oracle_geometry *g = G.d;
    float detect = interp_property(g, p.wavelength, G.surf(si, 0));
    float reflect_diffuse = interp_property(g, p.wavelength, G.surf(si, 3));
    float n2_eta = interp_property(g, p.wavelength, G.surf(si, 5));
    float n2_k = interp_property(g, p.wavelength, G.surf(si, 6));
    bool transmissive = g->surface_transmissive[si] != 0;
    float thickness = g->surface_thickness[si];

    cx n1(s.refractive_index1, 0.0f), n2(n2_eta, n2_k), n3(s.refractive_index2, 0.0f);
    const cx one(1.0f, 0.0f), two(2.0f, 0.0f);
    float cos_t1 = dot(p.direction, s.surface_normal);
    if (cos_t1 < 0.0f) cos_t1 = -cos_t1;
    float theta = acosf(cos_t1);
    cx cos1(cosf(theta), 0.0f), sin1(sinf(theta), 0.0f);
    float e = 2.0f * PI_F * thickness * 1.0e6 / p.wavelength;
    cx ratio13sin = cmul(cmul(cdiv(n1, n3), cdiv(n1, n3)), cmul(sin1, sin1));
    cx cos3 = csqrt_(one - ratio13sin);
    cx ratio12sin = cmul(cmul(cdiv(n1, n2), cdiv(n1, n2)), cmul(sin1, sin1));
    cx cos2 = csqrt_(one - ratio12sin);
    float u = cmul(n2, cos2).real(), v = cmul(n2, cos2).imag();

    float s_r, s_t, p_r, p_t, n_r, n_t;
    {
        cx n1c1 = cmul(n1, cos1), n2c2 = cmul(n2, cos2), n3c3 = cmul(n3, cos3);
        thin_film(cdiv(n1c1 - n2c2, n1c1 + n2c2), cdiv(n2c2 - n3c3, n2c2 + n3c3),
                  cdiv(cmul(two, n1c1), n1c1 + n2c2), cdiv(cmul(two, n2c2), n2c2 + n3c3),
                  cdiv(n3c3, n1c1), u, v, e, s_r, s_t);
    }
    {
        cx n2c1 = cmul(n2, cos1), n3c2 = cmul(n3, cos2), n2c3 = cmul(n2, cos3), n1c2 = cmul(n1, cos2);
        thin_film(cdiv(n2c1 - n1c2, n2c1 + n1c2), cdiv(n3c2 - n2c3, n3c2 + n2c3),
                  cdiv(cmul(cmul(two, n1), cos1), n2c1 + n1c2), cdiv(cmul(cmul(two, n2), cos2), n3c2 + n2c3),
                  cdiv(cmul(n3, cos3), cmul(n1, cos1)), u, v, e, p_r, p_t);
    }
    thin_film(cdiv(n1 - n2, n1 + n2), cdiv(n2 - n3, n2 + n3), cdiv(cmul(two, n1), n1 + n2),
              cdiv(cmul(two, n2), n2 + n3), cdiv(n3, n1), n2_eta, n2_k, e, n_r, n_t);

    float incident_angle = get_theta(s.surface_normal, -p.direction);
    float refracted_angle = asinf(sinf(incident_angle) * s.refractive_index1 / s.refractive_index2);
    f3 incident_plane_normal = cross(p.direction, s.surface_normal);
    float incident_plane_normal_length = norm(incident_plane_normal);
    if (incident_plane_normal_length < 1e-6f)
        incident_plane_normal = p.polarization;
    else
        incident_plane_normal = incident_plane_normal / incident_plane_normal_length;
    float normal_coefficient = dot(p.polarization, incident_plane_normal);
    float normal_probability = normal_coefficient * normal_coefficient;

    float transmit = normal_probability * s_t + (1.0f - normal_probability) * p_t;
    float transmit_normal_incidence = n_t;
    if (!transmissive) {
        transmit = 0.0f;
        transmit_normal_incidence = 0.0f;
    }
    float reflect = normal_probability * s_r + (1.0f - normal_probability) * p_r;
    float reflect_normal_incidence = n_r;
    float absorb = 1.0f - transmit - reflect;
    float absorb_normal_incidence = 1.0f - transmit_normal_incidence - reflect_normal_incidence;
    detect /= absorb_normal_incidence;
    if (use_weights && p.weight > WEIGHT_LOWER_THRESHOLD && absorb < (1.0f - WEIGHT_LOWER_THRESHOLD)) {
        float survive = 1.0f - absorb;
        absorb = 0.0f;
        p.weight *= survive;
        detect /= survive;
        reflect /= survive;
        transmit /= survive;
    }
    if (use_weights && detect > 0.0f) {
        p.history |= SURFACE_DETECT;
        p.weight *= detect;
        return BREAK;
    }
    float uniform_sample = curand_uniform(&rng);
    if (uniform_sample < absorb) {
        float uniform_sample_detect = curand_uniform(&rng);
        if (uniform_sample_detect < detect) p.history |= SURFACE_DETECT;
        else p.history |= SURFACE_ABSORB;
        return BREAK;
    } else if (uniform_sample < absorb + reflect || !transmissive) {
        float uniform_sample_reflect = curand_uniform(&rng);
        if (uniform_sample_reflect < reflect_diffuse) return propagate_at_diffuse_reflector(p, s, rng);
        else return propagate_at_specular_reflector(p, s);
    } else {
        p.direction = rotate(s.surface_normal, PI_F - refracted_angle, incident_plane_normal);
        p.polarization = cross(incident_plane_normal, p.direction);
        p.polarization = p.polarization / norm(p.polarization);
        p.history |= SURFACE_TRANSMIT;
        return CONTINUE;
    }
}

// photon.h:592-637
int propagate_at_wls(Photon &p, State &s, Rng &rng, const Geo &G, int si, bool use_weights) {
    const oracle_geometry *g = G.d;
    float absorb = interp_property(g, p.wavelength, G.surf(si, 1));
    float reflect_specular = interp_property(g, p.wavelength, G.surf(si, 4));
    float reflect_diffuse = interp_property(g, p.wavelength, G.surf(si, 3));
    float reemit = interp_property(g, p.wavelength, G.surf(si, 2));
    float uniform_sample = curand_uniform(&rng);
    if (use_weights && p.weight > WEIGHT_LOWER_THRESHOLD && absorb < (1.0f - WEIGHT_LOWER_THRESHOLD)) {
        float survive = 1.0f - absorb;
        absorb = 0.0f;
        p.weight *= survive;
        reflect_diffuse /= survive;
        reflect_specular /= survive;
    }
    if (uniform_sample < absorb) {
        float uniform_sample_reemit = curand_uniform(&rng);
        if (uniform_sample_reemit < reemit) {
            p.history |= SURFACE_REEMIT;
            p.wavelength = sample_cdf(&rng, (int)g->nwavelengths, g->wavelength_start, g->wavelength_step, G.surf(si, 7));
            p.direction = uniform_sphere(&rng);
            p.polarization = cross(uniform_sphere(&rng), p.direction);
            p.polarization = p.polarization / norm(p.polarization);
            return CONTINUE;
        } else {
            p.history |= SURFACE_ABSORB;
            return BREAK;
        }
    } else if (uniform_sample < absorb + reflect_specular + reflect_diffuse) {
        float uniform_sample_reflect = curand_uniform(&rng) * (reflect_specular + reflect_diffuse);
        if (uniform_sample_reflect < reflect_specular) return propagate_at_specular_reflector(p, s);
        else return propagate_at_diffuse_reflector(p, s, rng);
    } else {
        p.history |= SURFACE_TRANSMIT;
        return PASS;
    }
}

// photon.h:640-670
int propagate_at_dichroic(Photon &p, State &s, Rng &rng, const Geo &G, int si) {
    const oracle_geometry *g = G.d;
    float incident_angle = get_theta(s.surface_normal, -p.direction);
    unsigned nangles = g->dichroic_nangles[si];
    const size_t nw = g->nwavelengths;
    const float *angles = g->dichroic_angles + G.dich_first[si];
    const float *refl = g->dichroic_reflect + G.dich_first[si] * nw;
    const float *tran = g->dichroic_transmit + G.dich_first[si] * nw;
    float idx = interp_idx(incident_angle, nangles, angles);
    unsigned int iidx = (int)idx;
    unsigned int iidx_hi = iidx < nangles - 2 ? iidx + 1 : iidx;
    float reflect_prob_low = interp_property(g, p.wavelength, refl + iidx * nw);
    float reflect_prob_high = interp_property(g, p.wavelength, refl + iidx_hi * nw);
    float transmit_prob_low = interp_property(g, p.wavelength, tran + iidx * nw);
    float transmit_prob_high = interp_property(g, p.wavelength, tran + iidx_hi * nw);
    float reflect_prob = reflect_prob_low + (reflect_prob_high - reflect_prob_low) * (idx - iidx);
    float transmit_prob = transmit_prob_low + (transmit_prob_high - transmit_prob_low) * (idx - iidx);
    float uniform_sample = curand_uniform(&rng);
    if (uniform_sample < reflect_prob) {
        return propagate_at_specular_reflector(p, s);
    } else if (uniform_sample < transmit_prob + reflect_prob) {
        p.history |= SURFACE_TRANSMIT;
        return PASS;
    } else {
        p.history |= SURFACE_ABSORB;
        return BREAK;
    }
}

// photon.h:672-733
int propagate_at_surface(Photon &p, State &s, Rng &rng, const Geo &G, bool use_weights) {
    const oracle_geometry *g = G.d;
    int si = s.surface_index;
    unsigned model = g->surface_model[si];
    if (model == SURFACE_COMPLEX) return propagate_complex(p, s, rng, G, si, use_weights);
    else if (model == SURFACE_WLS) return propagate_at_wls(p, s, rng, G, si, use_weights);
    else if (model == SURFACE_DICHROIC) return propagate_at_dichroic(p, s, rng, G, si);
    float detect = interp_property(g, p.wavelength, G.surf(si, 0));
    float absorb = interp_property(g, p.wavelength, G.surf(si, 1));
    float reflect_diffuse = interp_property(g, p.wavelength, G.surf(si, 3));
    float reflect_specular = interp_property(g, p.wavelength, G.surf(si, 4));
    float uniform_sample = curand_uniform(&rng);
    if (use_weights && p.weight > WEIGHT_LOWER_THRESHOLD && absorb < (1.0f - WEIGHT_LOWER_THRESHOLD)) {
        float survive = 1.0f - absorb;
        absorb = 0.0f;
        p.weight *= survive;
        detect /= survive;
        reflect_diffuse /= survive;
        reflect_specular /= survive;
    }
    if (use_weights && detect > 0.0f) {
        p.history |= SURFACE_DETECT;
        p.weight *= detect;
        return BREAK;
    }
    if (uniform_sample < absorb) {
        p.history |= SURFACE_ABSORB;
        return BREAK;
    } else if (uniform_sample < absorb + detect) {
        p.history |= SURFACE_DETECT;
        return BREAK;
    } else if (uniform_sample < absorb + detect + reflect_diffuse)
        return propagate_at_diffuse_reflector(p, s, rng);
    else if (uniform_sample < absorb + detect + reflect_diffuse + reflect_specular)
        return propagate_at_specular_reflector(p, s);
    else
        return PASS;
}

}  // namespace

// propagate.cu:217-319 (kernel body) under the host loop of gpu/photon.py:193-259: the net
// effect of the launch/queue machinery is "every non-terminal photon takes up to max_steps steps"
extern "C" void oracle_propagate(const oracle_geometry *g, uint64_t n, float *pos, float *dir, float *pol,
                                 float *wavelengths, float *t, int32_t *last_hit, uint32_t *flags,
                                 float *weights, uint64_t seed, uint32_t id_base, uint32_t epoch,
                                 int max_steps, int use_weights, int scatter_first_arg, uint64_t *out_steps) {
    Geo G(g);
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t id = 0; id < (int64_t)n; id++) {
        Photon p;
        p.position = mk(pos[3 * id], pos[3 * id + 1], pos[3 * id + 2]);
        p.direction = mk(dir[3 * id], dir[3 * id + 1], dir[3 * id + 2]);
        p.direction = p.direction / norm(p.direction);
        p.polarization = mk(pol[3 * id], pol[3 * id + 1], pol[3 * id + 2]);
        p.polarization = p.polarization / norm(p.polarization);
        p.wavelength = wavelengths[id];
        p.time = t[id];
        p.last_hit_triangle = last_hit[id];
        p.history = flags[id];
        p.weight = weights[id];
        unsigned count = 0, tri_count = 0;
        int steps = 0;
        if (!(p.history & TERMINAL)) {
            Rng rng;
            rng_init(rng, seed, id_base + (uint32_t)id, epoch, 0u);
            int scatter_first = scatter_first_arg;
            State s;
            while (steps < max_steps) {
                steps++;
                int command;
                if (std::isnan(p.direction.x * p.direction.y * p.direction.z * p.position.x * p.position.y * p.position.z)) {
                    p.history |= NO_HIT | NAN_ABORT;
                    break;
                }
                fill_state(s, p, G, count, tri_count);
                if (p.last_hit_triangle == -1) break;
                command = propagate_to_boundary(p, s, rng, G, use_weights != 0, scatter_first);
                scatter_first = 0;
                if (command == BREAK) break;
                if (command == CONTINUE) continue;
                if (s.surface_index != -1) {
                    command = propagate_at_surface(p, s, rng, G, use_weights != 0);
                    if (command == BREAK) break;
                    if (command == CONTINUE) continue;
                }
                propagate_at_boundary(p, s, rng);
            }
            pos[3 * id] = p.position.x; pos[3 * id + 1] = p.position.y; pos[3 * id + 2] = p.position.z;
            dir[3 * id] = p.direction.x; dir[3 * id + 1] = p.direction.y; dir[3 * id + 2] = p.direction.z;
            pol[3 * id] = p.polarization.x; pol[3 * id + 1] = p.polarization.y; pol[3 * id + 2] = p.polarization.z;
            wavelengths[id] = p.wavelength;
            t[id] = p.time;
            flags[id] = p.history;
            last_hit[id] = p.last_hit_triangle;
            weights[id] = p.weight;
        }
        if (out_steps) {
            out_steps[3 * id] = (uint64_t)steps;
            out_steps[3 * id + 1] = count;
            out_steps[3 * id + 2] = tri_count;
        }
    }
}

// daq.cu:35-86 (run_daq) between gpu/daq.py:54-58 (begin_acquire) and :93-99 (end_acquire)
extern "C" void oracle_run_daq(const oracle_geometry *g, uint64_t n, const float *t, const uint32_t *flags,
                               const int32_t *last_hit, const float *weights, uint64_t seed, uint32_t id_base,
                               uint32_t epoch, float global_weight, float *out_t, float *out_q, uint32_t *out_hist) {
    const uint32_t nch = g->nchannels;
    std::vector<uint32_t> time_int(nch), q_int(nch, 0), hist(nch, 0);
    float maxtime = 1e9f;
    uint32_t maxtime_int;
    memcpy(&maxtime_int, &maxtime, 4);
    for (uint32_t i = 0; i < nch; i++) time_int[i] = maxtime_int;
    const float charge_unit = g->charge_cdf_x[g->charge_cdf_len - 1] / 65536.0f;  // gpu/detector.py:39
    for (uint64_t id = 0; id < n; id++) {  // serial: the atomics of the kernel commute
        int triangle_id = last_hit[id];
        if (triangle_id > -1) {
            int solid_id = g->solid_id[triangle_id];
            unsigned history = flags[id];
            int channel_index = g->solid_id_to_channel_index[solid_id];
            if (channel_index >= 0 && (history & SURFACE_DETECT)) {
                Rng rng;
                rng_init(rng, seed, id_base + (uint32_t)id, epoch, 1u);
                float weight = weights[id] * global_weight;
                if (curand_uniform(&rng) < weight) {
                    float time = t[id] + sample_cdf(&rng, (int)g->time_cdf_len, g->time_cdf_x, g->time_cdf_y);
                    uint32_t tint;
                    memcpy(&tint, &time, 4);  // float_to_sortable_int: raw bits (daq.cu:5-12)
                    float charge = sample_cdf(&rng, (int)g->charge_cdf_len, g->charge_cdf_x, g->charge_cdf_y);
                    unsigned charge_int = (unsigned)roundf(charge / charge_unit);
                    if (tint < time_int[channel_index]) time_int[channel_index] = tint;  // atomicMin
                    q_int[channel_index] += charge_int;                                  // atomicAdd
                    hist[channel_index] |= history;                                      // atomicOr
                }
            }
        }
    }
    for (uint32_t i = 0; i < nch; i++) {
        memcpy(&out_t[i], &time_int[i], 4);        // convert_sortable_int_to_float
        out_q[i] = q_int[i] * charge_unit;         // convert_charge_int_to_float
        out_hist[i] = hist[i];
    }
}

// daq.cu:88-150 (run_daq_many) between gpu/daq.py:54-58 and :93-99, ndaq replicas of every channel
// (replica k at channel_index + k * nchannels, gpu/daq.py:44-45 stride).  The reference runs one
// block per photon whose thread i serves replicas i, i + blockDim, ...; each (photon, replica) pair
// draws from its own stream here (stream word 0x10000 + k), as the engine does: accept draw, the
// two uniforms of curand_normal's Box-Muller, time CDF, charge CDF.
// Reference quirk kept as an option: convert_charge_int_to_float (daq.cu:161-172) converts only the
// first nchannels charges, so replicas >= 1 keep q = 0 there; `convert_all` != 0 converts them all
// (what the engine does, engine_daq.cu).
extern "C" void oracle_run_daq_many(const oracle_geometry *g, uint64_t n, const float *t, const uint32_t *flags,
                                    const int32_t *last_hit, const float *weights, uint64_t seed, uint32_t id_base,
                                    uint32_t epoch, float global_weight, int ndaq, int convert_all, float *out_t,
                                    float *out_q, uint32_t *out_hist) {
    const uint32_t nch = g->nchannels;
    const size_t tot = (size_t)nch * ndaq;
    std::vector<uint32_t> time_int(tot), q_int(tot, 0), hist(tot, 0);
    float maxtime = 1e9f;
    uint32_t maxtime_int;
    memcpy(&maxtime_int, &maxtime, 4);
    for (size_t i = 0; i < tot; i++) time_int[i] = maxtime_int;
    const float charge_unit = g->charge_cdf_x[g->charge_cdf_len - 1] / 65536.0f;
    for (uint64_t id = 0; id < n; id++) {
        int triangle_id = last_hit[id];
        if (triangle_id <= -1) continue;
        int solid_id = g->solid_id[triangle_id];
        unsigned history = flags[id];
        int channel_index = g->solid_id_to_channel_index[solid_id];
        if (channel_index < 0 || !(history & SURFACE_DETECT)) continue;
        float photon_time = t[id];
        float weight = weights[id] * global_weight;
        for (int k = 0; k < ndaq; k++) {
            size_t channel_offset = (size_t)channel_index + (size_t)k * nch;
            Rng rng;
            rng_init(rng, seed, id_base + (uint32_t)id, epoch, 0x10000u + (uint32_t)k);
            if (curand_uniform(&rng) < weight) {
                float u1 = curand_uniform(&rng), u2 = curand_uniform(&rng);
                float normal = sqrtf(-2.0f * logf(u1)) * sinf(2.0f * 3.141592653589793f * u2);
                float time = photon_time + normal +
                             sample_cdf(&rng, (int)g->time_cdf_len, g->time_cdf_x, g->time_cdf_y);
                uint32_t tint;
                memcpy(&tint, &time, 4);
                float charge = sample_cdf(&rng, (int)g->charge_cdf_len, g->charge_cdf_x, g->charge_cdf_y);
                unsigned charge_int = (unsigned)roundf(charge / charge_unit);
                if (tint < time_int[channel_offset]) time_int[channel_offset] = tint;
                q_int[channel_offset] += charge_int;
                hist[channel_offset] |= history;
            }
        }
    }
    for (size_t i = 0; i < tot; i++) {
        memcpy(&out_t[i], &time_int[i], 4);
        out_q[i] = (convert_all || i < nch) ? q_int[i] * charge_unit : 0.0f;
        out_hist[i] = hist[i];
    }
}

// pdf.cu:9-32
extern "C" void oracle_bin_hits(int nchannels, const float *channel_q, const float *channel_time, uint32_t *hitcount,
                                int tbins, float tmin, float tmax, int qbins, float qmin, float qmax, uint32_t *pdf) {
    for (int id = 0; id < nchannels; id++) {
        unsigned int q = (unsigned int)channel_q[id];
        float t = channel_time[id];
        if (t < 1e8 && t >= tmin && t < tmax && q >= qmin && q < qmax) {
            hitcount[id] += 1;
            int tbin = (t - tmin) / (tmax - tmin) * tbins;
            int qbin = (q - qmin) / (qmax - qmin) * qbins;
            size_t bin = (size_t)id * ((size_t)tbins * qbins) + (size_t)tbin * qbins + qbin;
            pdf[bin] += 1;
        }
    }
}

// pdf.cu:34-149 + sorting.h piksrt: the reference's two passes, kept separate — pass 1 fills a
// per-hit-channel work queue with the replicas whose distance must be sorted in, pass 2 appends
// them to the valid prefix of the channel's list, sorts and keeps the first min_bin_content.
extern "C" void oracle_accumulate_pdf_eval(int nchannels, int ndaq, const uint32_t *event_hit,
                                           const float *event_time, const float *mc_time, uint32_t *hitcount,
                                           uint32_t *bincount, float min_twidth, float tmin, float tmax,
                                           int min_bin_content, float *nearest_mc) {
    std::vector<int> queue;
    std::vector<float> table;
    int hit_id = -1;
    for (int ch = 0; ch < nchannels; ch++) {
        float ch_hitcount = hitcount[ch];   // float counters, pdf.cu:52-53
        float ch_bincount = bincount[ch];
        const float ev_time = event_time[ch];
        const int ev_hit = event_hit[ch];
        if (ev_hit) hit_id++;
        queue.clear();
        for (int i = 0; i < ndaq; i++) {
            int read_offset = nchannels * i + ch;
            float t = mc_time[read_offset];
            if (t >= 1e8) continue;
            if (t < tmin || t > tmax) continue;
            ch_hitcount += 1;
            if (!ev_hit) continue;
            float distance = fabsf(t - ev_time);
            if (distance < min_twidth / 2.0) ch_bincount += 1;
            if (ch_bincount < min_bin_content) queue.push_back(read_offset);
        }
        hitcount[ch] = ch_hitcount;
        bincount[ch] = ch_bincount;
        if (!ev_hit) continue;
        // pass 2 (pdf.cu:151-219) for this hit channel
        float *list = nearest_mc + (size_t)min_bin_content * hit_id;
        table.clear();
        for (int i = 0; i < min_bin_content; i++) {
            if (list[i] > 1e8) break;
            table.push_back(list[i]);
        }
        for (int off : queue) table.push_back(fabsf(mc_time[off] - ev_time));
        // sorting.h:18-34 piksrt sorts ascending; for NaN-free floats any ascending sort gives the same array
        std::sort(table.begin(), table.end());
        size_t keep = table.size() < (size_t)min_bin_content ? table.size() : (size_t)min_bin_content;
        for (size_t i = 0; i < keep; i++) list[i] = table[i];
    }
}

// pdf.cu:223-266.  nvcc (default -fmad=true, which the reference build keeps) contracts
// `m += x * x` into one fused multiply-add; fmaf states that rounding explicitly.
extern "C" void oracle_accumulate_moments(int time_only, int nchannels, const float *mc_time,
                                          const float *mc_charge, float tmin, float tmax, float qmin, float qmax,
                                          uint32_t *mom0, float *t_mom1, float *t_mom2, float *q_mom1,
                                          float *q_mom2) {
    for (int id = 0; id < nchannels; id++) {
        float t = mc_time[id];
        if (time_only) {
            if (t < tmin || t > tmax) continue;
            mom0[id] += 1;
            t_mom1[id] += t;
            t_mom2[id] = fmaf(t, t, t_mom2[id]);
        } else {
            float q = mc_charge[id];
            if (t < tmin || t > tmax || q < qmin || q > qmax) continue;
            mom0[id] += 1;
            t_mom1[id] += t;
            t_mom2[id] = fmaf(t, t, t_mom2[id]);
            q_mom1[id] += q;
            q_mom2[id] = fmaf(q, q, q_mom2[id]);
        }
    }
}

// pdf.cu:268-368.  The device code runs exp/erf/divide in fast-math float; this restatement
// evaluates each new term in double and rounds it once, so it is the tolerance anchor, not a
// bit pattern.
extern "C" void oracle_accumulate_kernel_eval(int time_only, int nchannels, const uint32_t *event_hit,
                                              const float *event_time, const float *event_charge,
                                              const float *mc_time, const float *mc_charge, float tmin,
                                              float tmax, float qmin, float qmax, const float *inv_tbw,
                                              const float *inv_qbw, uint32_t *hitcount, float *time_pdf,
                                              float *charge_pdf) {
    const double invroot2 = 0.70710678118654746, rootPiBy2 = 1.2533141373155001;
    auto norm_of = [&](double lo, double hi, double x, double inv_bw) {
        double norm = hi - lo;
        if (inv_bw > 0.0) norm = (erf((hi - x) * inv_bw * invroot2) - erf((lo - x) * inv_bw * invroot2)) * rootPiBy2;
        return norm;
    };
    for (int id = 0; id < nchannels; id++) {
        float t = mc_time[id];
        if (time_only) {
            if (t < tmin || t > tmax) continue;
            hitcount[id] += 1;
            if (!event_hit[id]) continue;
            double ibw = inv_tbw[id];
            double arg = ((double)t - event_time[id]) * ibw;
            double term = exp(-0.5 * arg * arg) * ibw;   // pdf.cu:303: time-only multiplies by 1/bw
            time_pdf[id] += (float)(term / norm_of(tmin, tmax, t, ibw));
        } else {
            float q = mc_charge[id];
            if (t < tmin || t > tmax || q < qmin || q > qmax) continue;
            hitcount[id] += 1;
            if (!event_hit[id]) continue;
            double ibw = inv_tbw[id];
            double arg = ((double)t - event_time[id]) * ibw;
            time_pdf[id] += (float)(exp(-0.5 * arg * arg) / norm_of(tmin, tmax, t, ibw));
            ibw = inv_qbw[id];
            arg = ((double)q - event_charge[id]) * ibw;
            charge_pdf[id] += (float)(exp(-0.5 * arg * arg) / norm_of(qmin, qmax, q, ibw));
        }
    }
}

// render.cu:11-33 get_color
static inline void get_color(const f3 &direction, const Triangle &t, unsigned rgba, float out[4]) {
    f3 v01 = t.v1 - t.v0, v12 = t.v2 - t.v1;
    f3 c = cross_fused(v01, v12);
    float len = sqrtf(dot_fused(c, c));
    f3 sn = mk(c.x / len, c.y / len, c.z / len);
    float cos_theta = dot_fused(sn, mk(-direction.x, -direction.y, -direction.z));
    if (cos_theta < 0.0f) cos_theta = -cos_theta;
    unsigned a0 = 0xff & (rgba >> 24), r0 = 0xff & (rgba >> 16), g0 = 0xff & (rgba >> 8), b0 = 0xff & rgba;
    out[0] = r0 * cos_theta; out[1] = g0 * cos_theta; out[2] = b0 * cos_theta;
    out[3] = (255 - a0) / 255.0f;
}

// render.cu:35-181 render.  dx [n][alpha_depth], color [n][alpha_depth][4], dxlen [n] are in/out
// (keep_last_render = the caller does not zero dxlen).  The device code evaluates the colour and the
// blend with fast-math division / rsqrt, so pixels can differ from this restatement in the last
// unit of a channel; the hit lists (dxlen, dx) are exact.
extern "C" void oracle_render(const oracle_geometry *g, const uint32_t *colors, uint64_t nrays, const float *pos,
                              const float *dir, uint32_t alpha_depth, uint32_t *pixels, float *dx_all,
                              uint32_t *dxlen, float *color_all, uint32_t bg_color) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t id = 0; id < (int64_t)nrays; id++) {
        f3 origin = mk(pos[3 * id], pos[3 * id + 1], pos[3 * id + 2]);
        f3 direction = mk(dir[3 * id], dir[3 * id + 1], dir[3 * id + 2]);
        unsigned n = dxlen[id];
        float distance;
        Node root = get_node(g, 0);
        f3 inv = mk(1.0f / direction.x, 1.0f / direction.y, 1.0f / direction.z);
        f3 nod = mk(-origin.x * inv.x, -origin.y * inv.y, -origin.z * inv.z);
        if (n < 1 && !intersect_node(nod, inv, root, -1.0f)) {
            pixels[id] = bg_color;
            continue;
        }
        std::vector<unsigned> child_ptr_stack(STACK_SIZE), nchild_ptr_stack(STACK_SIZE);
        child_ptr_stack[0] = root.child;
        nchild_ptr_stack[0] = root.nchild;
        int curr = 0;
        float *dx = dx_all + (size_t)id * alpha_depth;
        float *color_a = color_all + (size_t)id * alpha_depth * 4;
        while (curr >= 0) {
            unsigned first_child = child_ptr_stack[curr], nchild = nchild_ptr_stack[curr];
            curr--;
            for (unsigned i = first_child; i < first_child + nchild; i++) {
                Node node = get_node(g, i);
                if (!intersect_node(nod, inv, node, -1.0f)) continue;
                if (node.nchild == 0) {
                    Triangle t = get_triangle(g, node.child);
                    if (intersect_triangle(origin, direction, t, distance)) {
                        float col[4];
                        if (n < 1) {
                            dx[0] = distance;
                            get_color(direction, t, colors[node.child], col);
                            memcpy(color_a, col, 16);
                        } else {
                            // sorting.h:62-98 searchsorted on an ascending table = first index whose entry
                            // is >= distance; insert() shifts the rest back and drops the last
                            unsigned long j = std::lower_bound(dx, dx + n, distance) - dx;
                            if (j <= alpha_depth - 1) {
                                for (unsigned long m = alpha_depth - 1; m > j; m--) {
                                    dx[m] = dx[m - 1];
                                    memcpy(color_a + 4 * m, color_a + 4 * (m - 1), 16);
                                }
                                dx[j] = distance;
                                get_color(direction, t, colors[node.child], col);
                                memcpy(color_a + 4 * j, col, 16);
                            }
                        }
                        if (n < alpha_depth) n++;
                    }
                } else if (curr < STACK_SIZE - 1) {
                    curr++;
                    child_ptr_stack[curr] = node.child;
                    nchild_ptr_stack[curr] = node.nchild;
                }
            }
        }
        if (n < 1) {
            pixels[id] = bg_color;
            continue;
        }
        dxlen[id] = n;
        float scale = 1.0f, fr = 0.0f, fg = 0.0f, fb = 0.0f;
        for (unsigned i = 0; i < n; i++) {
            float alpha = color_a[4 * i + 3];
            fr += scale * color_a[4 * i] * alpha;
            fg += scale * color_a[4 * i + 1] * alpha;
            fb += scale * color_a[4 * i + 2] * alpha;
            scale *= (1.0f - alpha);
        }
        float alpha = ((float)((bg_color & 0xFF000000) >> 24)) / 255.0;
        fr += scale * ((float)((bg_color & 0xFF0000) >> 16)) * alpha;
        fg += scale * ((float)((bg_color & 0xFF00) >> 8)) * alpha;
        fb += scale * ((float)(bg_color & 0xFF)) * alpha;
        scale *= (1.0f - alpha);
        unsigned a = n < alpha_depth ? (unsigned)floorf(255 * (1.0f - scale)) : 255u;
        unsigned red = (unsigned)floorf(fr / (1.0f - scale));
        unsigned green = (unsigned)floorf(fg / (1.0f - scale));
        unsigned blue = (unsigned)floorf(fb / (1.0f - scale));
        pixels[id] = a << 24 | red << 16 | green << 8 | blue;
    }
}

// =======================================================================================
// BVH construction (bvh.cu, bvh/grid.py, gpu/bvh.py)
// =======================================================================================
namespace {
// bvh.cu:42-52
inline unsigned long long spread3_16(unsigned int input) {
    unsigned long long x = input;
    x = (x | (x << 16)) & 0x00000000FF0000FFull;
    x = (x | (x << 8)) & 0x000000F00F00F00Full;
    x = (x | (x << 4)) & 0x00000C30C30C30C3ull;
    x = (x | (x << 2)) & 0x0000249249249249ull;
    return x;
}
// bvh.cu:65-69 (truncating)
inline unsigned quantize(float v, float world_origin, float world_scale) {
    return (unsigned int)((v - world_origin) / world_scale);
}
struct U4 { uint32_t x, y, z, w; };
}  // namespace

// bvh.cu:148-203
extern "C" void oracle_make_leaves(const float *vertices, const uint32_t *triangles, uint64_t ntriangles,
                                   const float wo[3], float ws, uint32_t *leaf_nodes, uint64_t *morton_codes) {
#pragma omp parallel for
    for (int64_t tid = 0; tid < (int64_t)ntriangles; tid++) {
        const uint32_t *tr = triangles + 3 * tid;
        const float *a = vertices + 3 * (size_t)tr[0], *b = vertices + 3 * (size_t)tr[1], *c = vertices + 3 * (size_t)tr[2];
        float lo[3], hi[3], ce[3];
        for (int k = 0; k < 3; k++) {
            lo[k] = fminf(fminf(a[k], b[k]), c[k]);
            hi[k] = fmaxf(fmaxf(a[k], b[k]), c[k]);
            ce[k] = ((a[k] + b[k]) + c[k]) / 3.0f;
        }
        unsigned ql[3], qu[3], qc[3];
        for (int k = 0; k < 3; k++) {
            ql[k] = quantize(lo[k], wo[k], ws);
            if (ql[k] > 0) ql[k]--;
            qu[k] = quantize(hi[k], wo[k], ws) + 1;
            qc[k] = quantize(ce[k], wo[k], ws);
        }
        unsigned long long morton = spread3_16(qc[0]) | (spread3_16(qc[1]) << 1) | (spread3_16(qc[2]) << 2);
        uint32_t *out = leaf_nodes + 4 * tid;
        out[0] = ql[0] | (qu[0] << 16);
        out[1] = ql[1] | (qu[1] << 16);
        out[2] = ql[2] | (qu[2] << 16);
        out[3] = (uint32_t)tid;
        morton_codes[tid] = morton;
    }
}

// bvh/grid.py:11-95 make_recursive_grid_bvh (+ gpu/bvh.py:17-130,239-267)
extern "C" uint64_t oracle_build_bvh_from_leaves(const uint32_t *leaves_in, const uint64_t *codes_in,
                                                 uint64_t ntriangles, int target_degree,
                                                 uint32_t *nodes_out, uint64_t *layer_offsets,
                                                 uint32_t *nlayers);

extern "C" uint64_t oracle_build_bvh(const float *vertices, uint64_t nvertices, const uint32_t *triangles,
                                     uint64_t ntriangles, int target_degree, uint32_t *nodes_out,
                                     uint64_t *layer_offsets, uint32_t *nlayers, float world_origin[3],
                                     float *world_scale) {
    // world coordinates: gpu/bvh.py:43-48
    float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (uint64_t i = 0; i < nvertices; i++)
        for (int k = 0; k < 3; k++) {
            mn[k] = fminf(mn[k], vertices[3 * i + k]);
            mx[k] = fmaxf(mx[k], vertices[3 * i + k]);
        }
    float ext = fmaxf(fmaxf(mx[0] - mn[0], mx[1] - mn[1]), mx[2] - mn[2]);
    float ws = ext / 65534.0f;
    if (!(ws > 0.f)) ws = 1.0f;
    for (int k = 0; k < 3; k++) world_origin[k] = mn[k];
    *world_scale = ws;

    std::vector<U4> leaves(ntriangles);
    std::vector<uint64_t> codes(ntriangles);
    oracle_make_leaves(vertices, triangles, ntriangles, mn, ws, (uint32_t *)leaves.data(), codes.data());
    return oracle_build_bvh_from_leaves((const uint32_t *)leaves.data(), codes.data(), ntriangles,
                                        target_degree, nodes_out, layer_offsets, nlayers);
}

// the host part of make_recursive_grid_bvh (bvh/grid.py:24-95) on given leaves + Morton codes
// (the leaves of the reference come from its GPU kernel, whose approximate division the CPU
// cannot reproduce bit for bit; tests feed the device-made leaves through here)
extern "C" uint64_t oracle_build_bvh_from_leaves(const uint32_t *leaves_in, const uint64_t *codes_in,
                                                 uint64_t ntriangles, int target_degree,
                                                 uint32_t *nodes_out, uint64_t *layer_offsets,
                                                 uint32_t *nlayers) {
    const unsigned MAX_CHILD = 15;  // 2**(32-CHILD_BITS) - 1
    const U4 *leaves = (const U4 *)leaves_in;
    const uint64_t *codes = codes_in;
    // rearrange in morton order (argsort; stable here, unspecified in numpy)
    std::vector<uint32_t> order(ntriangles);
    std::iota(order.begin(), order.end(), 0u);
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return codes[a] < codes[b]; });
    std::vector<U4> top(ntriangles);
    std::vector<uint64_t> morton(ntriangles);
    for (uint64_t i = 0; i < ntriangles; i++) { top[i] = leaves[order[i]]; morton[i] = codes[order[i]]; }

    std::vector<std::vector<U4>> layers;  // bottom-up
    layers.push_back(top);
    auto count_unique = [](const std::vector<uint64_t> &a) {
        uint64_t c = 1;
        for (size_t i = 1; i < a.size(); i++) c += a[i] != a[i - 1];
        return c;
    };
    while (layers.back().size() > 1) {
        const std::vector<U4> &cur = layers.back();
        uint64_t nnodes = cur.size();
        uint64_t nunique = count_unique(morton);
        while ((double)nnodes / (double)nunique < (double)target_degree && nunique > 1) {
            for (auto &m : morton) m >>= 1;
            nunique = count_unique(morton);
        }
        // group equal codes; split groups with more than MAX_CHILD children (grid.py:43-71)
        std::vector<uint32_t> first_child, nchild;
        std::vector<uint64_t> parent_codes;
        uint64_t i = 0;
        while (i < nnodes) {
            uint64_t j = i + 1;
            while (j < nnodes && morton[j] == morton[i]) j++;
            for (uint64_t f = i; f < j; f += MAX_CHILD) {
                first_child.push_back((uint32_t)f);
                nchild.push_back((uint32_t)std::min<uint64_t>(MAX_CHILD, j - f));
                parent_codes.push_back(morton[i]);
            }
            i = j;
        }
        // make_parents_detailed: bvh.cu:269-308
        std::vector<U4> parents(first_child.size());
        for (size_t p = 0; p < parents.size(); p++) {
            U4 c0 = cur[first_child[p]];
            uint32_t lo[3] = {c0.x & 0xFFFF, c0.y & 0xFFFF, c0.z & 0xFFFF}, hi[3] = {c0.x >> 16, c0.y >> 16, c0.z >> 16};
            for (uint32_t k = 1; k < nchild[p]; k++) {
                U4 c = cur[first_child[p] + k];
                lo[0] = std::min(lo[0], c.x & 0xFFFF); lo[1] = std::min(lo[1], c.y & 0xFFFF); lo[2] = std::min(lo[2], c.z & 0xFFFF);
                hi[0] = std::max(hi[0], c.x >> 16); hi[1] = std::max(hi[1], c.y >> 16); hi[2] = std::max(hi[2], c.z >> 16);
            }
            parents[p] = U4{hi[0] << 16 | lo[0], hi[1] << 16 | lo[1], hi[2] << 16 | lo[2],
                            (nchild[p] << CHILD_BITS) | first_child[p]};
        }
        layers.push_back(parents);
        morton = parent_codes;
    }
    // concatenate_layers: gpu/bvh.py:239-267 + copy_and_offset bvh.cu:364-384
    uint64_t total = 0;
    for (auto &l : layers) total += l.size();
    if (total > 2 * ntriangles + 1024) return 0;
    U4 *nodes = (U4 *)nodes_out;
    std::vector<uint64_t> bounds;
    uint64_t pos = 0;
    bounds.push_back(0);
    for (int li = (int)layers.size() - 1; li >= 0; li--) {
        const std::vector<U4> &l = layers[li];
        uint64_t end = pos + l.size();
        uint32_t child_offset = (li == 0) ? 0u : (uint32_t)end;  // leaf layer keeps triangle ids
        for (size_t k = 0; k < l.size(); k++) {
            U4 v = l[k];
            uint32_t nc = v.w >> CHILD_BITS, ch = v.w & ~NCHILD_MASK;
            v.w = (nc << CHILD_BITS) | (ch + child_offset);
            nodes[pos + k] = v;
        }
        pos = end;
        bounds.push_back(pos);
    }
    // collapse_chains: gpu/bvh.py:114-130 + collapse_child bvh.cu:530-543 (deepest inner layer first)
    for (int li = (int)bounds.size() - 3; li >= 0; li--)
        for (uint64_t k = bounds[li]; k < bounds[li + 1]; k++) {
            U4 v = nodes[k];
            if ((v.w >> CHILD_BITS) == 1) nodes[k] = nodes[v.w & ~NCHILD_MASK];
        }
    *nlayers = (uint32_t)(bounds.size() - 1);
    for (size_t k = 0; k + 1 < bounds.size(); k++) layer_offsets[k] = bounds[k];
    return total;
}
