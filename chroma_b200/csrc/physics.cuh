// physics.cuh — per-photon optical physics (device code).
//
// Behavioural contract: chroma/cuda/photon.h (fill_state :83-135, rayleigh_scatter :167-191,
// propagate_to_boundary :193-308, propagate_at_boundary :310-363, reflectors :365-398,
// propagate_complex :400-590, propagate_at_wls :592-637, propagate_at_dichroic :640-670,
// propagate_at_surface :672-733), chroma/cuda/random.h:9-55 (samplers),
// chroma/cuda/geometry.h:62-75 (interp_property), chroma/cuda/interpolate.h,
// chroma/cuda/rotate.h:22-28.  The order of random draws follows the reference line by
// line; the generator itself is a counter-based Philox4x32-10 (the reference's XORWOW
// state is indexed by launch thread, chroma/cuda/propagate.cu:241, so only statistical
// agreement is defined for it).
#pragma once
#include <cuComplex.h>
#include "traverse.cuh"

#define CHROMA_SPEED_OF_LIGHT 299.792458f
#define CHROMA_PI 3.141592653589793f
#define WEIGHT_LOWER_THRESHOLD 0.0001f

enum { CMD_BREAK = 0, CMD_CONTINUE = 1, CMD_PASS = 2 };

// ---------------------------------------------------------------------------------------
// Philox4x32-10, counter = (photon id, block index, call epoch, stream), key = seed.
// ---------------------------------------------------------------------------------------
// Per-lane state is 6 registers; the key / fixed counter words are warp-uniform (RngKey).
struct RngKey {
    uint32_t k0, k1;   // key = seed
    uint32_t c2, c3;   // counter words 2, 3 = call epoch, stream
};
struct Rng {
    uint32_t id;          // counter word 0: photon id
    uint32_t blk;         // counter word 1: next block index
    uint32_t left;        // unread buffered words
    uint32_t b0, b1, b2;  // buffered words (next first)
};

__device__ __forceinline__ void philox_round(uint32_t &c0, uint32_t &c1, uint32_t &c2, uint32_t &c3,
                                             uint32_t k0, uint32_t k1) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
}

__device__ __forceinline__ RngKey rng_key(uint64_t seed, uint32_t epoch, uint32_t stream) {
    RngKey K;
    K.k0 = (uint32_t)seed; K.k1 = (uint32_t)(seed >> 32); K.c2 = epoch; K.c3 = stream;
    return K;
}

__device__ __forceinline__ void rng_init(Rng &g, uint32_t id) {
    g.id = id;
    g.blk = 0;
    g.left = 0;
    g.b0 = g.b1 = g.b2 = 0;
}

// One Philox4x32-10 block.  Out of line by default: inlined at every draw site the ten rounds were
// 3,200 of the propagate kernel's 8,800 instructions (141 KB of SASS), and in physics-heavy runs
// (config C4) 64 % of the warp-stall samples were instruction-cache misses (ncu "no_inst",
// profiles/r01_c4_*); a block is needed once per four draws, so the call costs little.
#ifndef RNG_NOINLINE
#define RNG_NOINLINE 1
#endif
#if RNG_NOINLINE
static __device__ __noinline__ uint4
#else
static __device__ __forceinline__ uint4
#endif
philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int i = 0; i < 10; i++) {
        philox_round(c0, c1, c2, c3, k0, k1);
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}

// -DRNG_DRAW_NOINLINE=1 (experiment for physics-bound runs, config C4): the buffer logic of a draw is
// ~40 instructions and was inlined at 35 draw sites (1,373 of the kernel's 5,920 instructions, nvdisasm
// line table); out of line, with the state passed and returned by value (registers), a draw site is a
// call and a few moves and all sites share one copy in the instruction cache.
#ifndef RNG_DRAW_NOINLINE
#define RNG_DRAW_NOINLINE 0
#endif
struct RngDraw { Rng g; uint32_t x; };
static __device__ __noinline__ RngDraw rng_u32_ool(Rng g, uint32_t k0, uint32_t k1, uint32_t c2, uint32_t c3) {
    RngDraw r;
    if (g.left == 0) {
        const uint4 b = philox4x32_10(g.id, g.blk, c2, c3, k0, k1);
        g.b0 = b.y; g.b1 = b.z; g.b2 = b.w;
        g.blk++;
        g.left = 3;
        r.x = b.x;
    } else {
        r.x = g.b0;
        g.b0 = g.b1; g.b1 = g.b2;
        g.left--;
    }
    r.g = g;
    return r;
}

__device__ __forceinline__ uint32_t rng_u32(Rng &g, const RngKey &K) {
#if RNG_DRAW_NOINLINE
    const RngDraw r = rng_u32_ool(g, K.k0, K.k1, K.c2, K.c3);
    g = r.g;
    return r.x;
#endif
    if (g.left == 0) {
        const uint4 r = philox4x32_10(g.id, g.blk, K.c2, K.c3, K.k0, K.k1);
        g.b0 = r.y; g.b1 = r.z; g.b2 = r.w;
        g.blk++;
        g.left = 3;
        return r.x;
    }
    uint32_t x = g.b0;
    g.b0 = g.b1; g.b1 = g.b2;
    g.left--;
    return x;
}

// number of 32-bit words consumed so far, and the inverse: reposition a stream there
// (the wavefront kernels keep only this count between steps)
__device__ __forceinline__ uint32_t rng_tell(const Rng &g) { return g.blk * 4u - g.left; }
__device__ __forceinline__ void rng_seek(Rng &g, const RngKey &K, uint32_t id, uint32_t ndrawn) {
    g.id = id;
    g.blk = ndrawn >> 2;
    g.left = 0;
    g.b0 = g.b1 = g.b2 = 0;
    uint32_t r = ndrawn & 3u;
    if (r) {
        (void)rng_u32(g, K);  // regenerates block ndrawn/4 and consumes word 0
        for (uint32_t i = 1; i < r; i++) (void)rng_u32(g, K);
    }
}

// curand_uniform: (0, 1]
__device__ __forceinline__ float rng_uniform(Rng &g, const RngKey &K) {
    return __fmaf_rn((float)rng_u32(g, K), 2.3283064365386963e-10f, 1.1641532182693481e-10f);
}

// random.h:9-13
__device__ __forceinline__ float uniform_range(Rng &g, const RngKey &K, float low, float high) {
    return low + rng_uniform(g, K) * (high - low);
}

// random.h:15-23
__device__ __forceinline__ float3 uniform_sphere(Rng &g, const RngKey &K) {
    float theta = uniform_range(g, K, 0.0f, 2 * CHROMA_PI);
    float u = uniform_range(g, K, -1.0f, 1.0f);
    float c = sqrtf(1.0f - u * u);
    return f3(c * cosf(theta), c * sinf(theta), u);
}

// interpolate.h:32-58 interp(x, n, xp, fp)
__device__ __forceinline__ float interp_tab(float x, int n, const float *xp, const float *fp) {
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

// interpolate.h:4-29 interp_idx
__device__ __forceinline__ float interp_idx(float x, int n, const float *xp) {
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

// random.h:27-31: sample_cdf with tabulated x
__device__ __forceinline__ float sample_cdf_tab(Rng &g, const RngKey &K, int n, const float *cdf_x,
                                                const float *cdf_y) {
    return interp_tab(rng_uniform(g, K), n, cdf_y, cdf_x);
}

// random.h:33-55: sample_cdf on a uniform grid
__device__ __forceinline__ float sample_cdf_uniform(Rng &g, const RngKey &K, int n, float x0, float delta,
                                                    const float *cdf_y) {
    float u = rng_uniform(g, K);
    int lower = 0, upper = n - 1;
    while (lower < upper - 1) {
        int half = (lower + upper) / 2;
        if (u < cdf_y[half]) upper = half; else lower = half;
    }
    float delta_cdf_y = cdf_y[upper] - cdf_y[lower];
    return x0 + delta * lower + delta * (u - cdf_y[lower]) / delta_cdf_y;
}

// ---------------------------------------------------------------------------------------
// optics tables
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ const float *opt_tab(const OpticsView &o, const float *s_opt,
                                                uint32_t off) {
    return (off + o.nw <= o.smem_floats) ? (s_opt + off) : (o.blob + off);
}

// geometry.h:62-75 interp_property
// out of line (24 call sites; see philox4x32_10 for why code size matters here)
static __device__ __noinline__ float interp_property_impl(float w0, float dw, uint32_t nw, float x, const float *fp) {
    if (x < w0) return fp[0];
    if (x > (w0 + (nw - 1) * dw)) return fp[nw - 1];
    int jl = (x - w0) / dw;
    return fp[jl] + (x - (w0 + jl * dw)) * (fp[jl + 1] - fp[jl]) / dw;
}
__device__ __forceinline__ float interp_property(const OpticsView &o, float x, const float *fp) {
    return interp_property_impl(o.w0, o.dw, o.nw, x, fp);
}

// ---------------------------------------------------------------------------------------
// photon + per-step state
// ---------------------------------------------------------------------------------------
struct Photon {
    float3 pos, dir, pol;
    float wavelength, time, weight;
    uint32_t history;
    int last_slot;  // leaf slot of last_hit_triangle (-1 none)
    int last_tri;   // triangle id
};

struct State {
    float3 normal;
    float n1, n2, absorption_length, scattering_length;
    int material1;      // material index (reference: Material *material1)
    int surface_index;  // -1 none
    float distance;
};

__device__ __forceinline__ float3 operator+(const float3 &a, const float3 &b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float3 operator-(const float3 &a, const float3 &b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ float3 operator-(const float3 &a) { return f3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float3 operator*(const float3 &a, float c) { return f3(a.x * c, a.y * c, a.z * c); }
__device__ __forceinline__ float3 operator*(float c, const float3 &a) { return f3(a.x * c, a.y * c, a.z * c); }
__device__ __forceinline__ float3 operator/(const float3 &a, float c) { return f3(a.x / c, a.y / c, a.z / c); }
__device__ __forceinline__ float dot3(const float3 &a, const float3 &b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float3 cross3(const float3 &a, const float3 &b) {
    return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ float norm3(const float3 &a) { return sqrtf(dot3(a, a)); }

// photon.h:68-75 convert: sign-extend the 8-bit material/surface index (0xFF -> -1)
__device__ __forceinline__ int convert8(int c) { return (c & 0x80) ? (int)(0xFFFFFF00u | c) : c; }

// photon.h:77-81
static __device__ __noinline__ float acos_clamped(float c) { return acosf(fmaxf(-1.0f, fminf(1.0f, c))); }
__device__ __forceinline__ float get_theta(const float3 &a, const float3 &b) { return acos_clamped(dot3(a, b)); }

// rotate.h:22-28
__device__ __forceinline__ float3 rotate3(const float3 &a, float phi, const float3 &n) {
    float cos_phi = cosf(phi), sin_phi = sinf(phi);
    return a * cos_phi + (n * dot3(a, n)) * (1.0f - cos_phi) + cross3(a, n) * sin_phi;
}

// ---------------------------------------------------------------------------------------
// photon.h:83-135 fill_state.  Returns false when nothing is hit (NO_HIT set).
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void fill_state_optics(State &s, const Photon &p, const OpticsView &o,
                                                  const float *s_opt, uint32_t code, float3 nrm);

// second half of fill_state: everything after intersect_mesh() returned (slot, dist)
__device__ __forceinline__ bool fill_state_post(State &s, Photon &p, const GeomView &g,
                                                const float *s_opt, int slot, float dist) {
    s.distance = dist;
    p.last_slot = slot;
    if (slot < 0) {
        p.last_tri = -1;
        p.history |= CHROMA_NO_HIT;
        return false;
    }
    TriRec t;
    load_tri(g, (uint32_t)slot, t);
    p.last_tri = (int)t.id;

    float3 v01 = t.v1 - t.v0;
    float3 v12 = t.v2 - t.v1;
    float3 nrm = cross3(v01, v12);
    nrm = nrm / norm3(nrm);
    fill_state_optics(s, p, g.opt, s_opt, t.code, nrm);
    return true;
}

// photon.h:96-133: material / surface indices of the triangle's code word, the side the photon comes
// from, and the four interpolated optical constants (`nrm`: unit normal of the triangle as stored)
__device__ __forceinline__ void fill_state_optics(State &s, const Photon &p, const OpticsView &o,
                                                  const float *s_opt, uint32_t code, float3 nrm) {
    int inner = convert8(0xFF & (code >> 24));
    int outer = convert8(0xFF & (code >> 16));
    s.surface_index = convert8(0xFF & (code >> 8));

    int m1, m2;
    if (dot3(nrm, -p.dir) > 0.0f) {  // outside to inside
        m1 = outer;
        m2 = inner;
    } else {                         // inside to outside
        m1 = inner;
        m2 = outer;
        nrm = -nrm;
    }
    s.normal = nrm;
    s.n1 = interp_property(o, p.wavelength, opt_tab(o, s_opt, o.mat_off + (m1 * 3 + 0) * o.nw));
    s.n2 = interp_property(o, p.wavelength, opt_tab(o, s_opt, o.mat_off + (m2 * 3 + 0) * o.nw));
    s.absorption_length =
        interp_property(o, p.wavelength, opt_tab(o, s_opt, o.mat_off + (m1 * 3 + 1) * o.nw));
    s.scattering_length =
        interp_property(o, p.wavelength, opt_tab(o, s_opt, o.mat_off + (m1 * 3 + 2) * o.nw));
    s.material1 = m1;
}

__device__ __forceinline__ bool fill_state(State &s, Photon &p, const GeomView &g,
                                           const uint4 *s_nodes, const float *s_opt,
                                           unsigned *stack, bool &overflow) {
    RayCtx r;
    ray_setup(r, p.pos, p.dir);
    float dist;
    int slot = intersect_mesh_slot<false>(g, s_nodes, r, p.last_slot, dist, stack, nullptr,
                                          nullptr, overflow);
    return fill_state_post(s, p, g, s_opt, slot, dist);
}

// photon.h:137-165 (SNOMAN rayscatter)
__device__ __forceinline__ float3 pick_new_direction(float3 axis, float theta, float phi) {
    float cos_theta, sin_theta, cos_phi, sin_phi;
    sincosf(theta, &sin_theta, &cos_theta);
    sincosf(phi, &sin_phi, &cos_phi);
    float sin_axis_theta = sqrtf(1.0f - axis.z * axis.z);
    float cos_axis_phi, sin_axis_phi;
    if (isnan(sin_axis_theta) || sin_axis_theta < 0.00001f) {
        cos_axis_phi = 1.0f;
        sin_axis_phi = 0.0f;
    } else {
        cos_axis_phi = axis.x / sin_axis_theta;
        sin_axis_phi = axis.y / sin_axis_theta;
    }
    float dirx = cos_theta * axis.x +
                 sin_theta * (axis.z * cos_phi * cos_axis_phi - sin_phi * sin_axis_phi);
    float diry = cos_theta * axis.y +
                 sin_theta * (cos_phi * axis.z * sin_axis_phi + sin_phi * cos_axis_phi);
    float dirz = cos_theta * axis.z - sin_theta * cos_phi * sin_axis_theta;
    return f3(dirx, diry, dirz);
}

// photon.h:167-191
__device__ __forceinline__ void rayleigh_scatter(Photon &p, Rng &rng, const RngKey &K) {
    float cos_theta =
        2.0f * cosf((acosf(1.0f - 2.0f * rng_uniform(rng, K)) - 2 * CHROMA_PI) / 3.0f);
    if (cos_theta > 1.0f) cos_theta = 1.0f;
    else if (cos_theta < -1.0f) cos_theta = -1.0f;
    float theta = acosf(cos_theta);
    float phi = uniform_range(rng, K, 0.0f, 2.0f * CHROMA_PI);
    p.dir = pick_new_direction(p.pol, theta, phi);
    if (1.0f - fabsf(cos_theta) < 1e-6f) {
        p.pol = pick_new_direction(p.pol, CHROMA_PI / 2.0f, phi);
    } else {
        p.pol = p.pol - cos_theta * p.dir;
    }
    p.dir = p.dir / norm3(p.dir);
    p.pol = p.pol / norm3(p.pol);
}

// photon.h:193-308
__device__ __forceinline__ int propagate_to_boundary(Photon &p, State &s, Rng &rng, const RngKey &K,
                                                     const GeomView &g, const float *s_opt,
                                                     bool use_weights, int scatter_first) {
    const OpticsView &o = g.opt;
    float absorption_distance = -s.absorption_length * logf(rng_uniform(rng, K));
    float scattering_distance = -s.scattering_length * logf(rng_uniform(rng, K));

    if (use_weights && p.weight > WEIGHT_LOWER_THRESHOLD)  // prevent absorption
        absorption_distance = 1e30f;
    else
        use_weights = false;

    if (scatter_first == 1) {  // force scatter
        float scatter_prob = 1.0f - expf(-s.distance / s.scattering_length);
        if (scatter_prob > WEIGHT_LOWER_THRESHOLD) {
            int i = 0;
            const int max_i = 1000;
            while (i < max_i && scattering_distance > s.distance) {
                scattering_distance = -s.scattering_length * logf(rng_uniform(rng, K));
                i++;
            }
            p.weight *= scatter_prob;
        }
    } else if (scatter_first == -1) {  // prevent scatter
        float no_scatter_prob = expf(-s.distance / s.scattering_length);
        if (no_scatter_prob > WEIGHT_LOWER_THRESHOLD) {
            int i = 0;
            const int max_i = 1000;
            while (i < max_i && scattering_distance <= s.distance) {
                scattering_distance = -s.scattering_length * logf(rng_uniform(rng, K));
                i++;
            }
            p.weight *= no_scatter_prob;
        }
    }

    if (absorption_distance <= scattering_distance) {
        if (absorption_distance <= s.distance) {
            p.time += absorption_distance / (CHROMA_SPEED_OF_LIGHT / s.n1);
            p.pos = p.pos + absorption_distance * p.dir;

            uint32_t ncomp = o.mat_ncomp[s.material1];
            if (ncomp == 0) {
                p.last_slot = -1; p.last_tri = -1;
                p.history |= CHROMA_BULK_ABSORB;
                return CMD_BREAK;
            }
            uint32_t cfirst = o.mat_comp_first[s.material1];
            float uniform_sample_comp = rng_uniform(rng, K);
            float prob = 0.0f;
            uint32_t comp;
            for (comp = 0;; comp++) {
                float comp_abs = interp_property(
                    o, p.wavelength, opt_tab(o, s_opt, o.comp_off + ((cfirst + comp) * 3 + 2) * o.nw));
                prob += s.absorption_length / comp_abs;
                if (uniform_sample_comp < prob || comp + 1 == ncomp) break;
            }
            float uniform_sample_reemit = rng_uniform(rng, K);
            float comp_reemit_prob = interp_property(
                o, p.wavelength, opt_tab(o, s_opt, o.comp_off + ((cfirst + comp) * 3 + 0) * o.nw));
            if (uniform_sample_reemit < comp_reemit_prob) {
                p.wavelength = sample_cdf_uniform(
                    rng, K, o.nw, o.w0, o.dw,
                    opt_tab(o, s_opt, o.comp_off + ((cfirst + comp) * 3 + 1) * o.nw));
                p.time += sample_cdf_uniform(rng, K, o.nt, o.t0, o.dt,
                                             o.blob + o.tcdf_off + (size_t)(cfirst + comp) * o.nt);
                p.dir = uniform_sphere(rng, K);
                p.pol = cross3(uniform_sphere(rng, K), p.dir);
                p.pol = p.pol / norm3(p.pol);
                p.last_slot = -1; p.last_tri = -1;
                p.history |= CHROMA_BULK_REEMIT;
                return CMD_CONTINUE;
            } else {
                p.last_slot = -1; p.last_tri = -1;
                p.history |= CHROMA_BULK_ABSORB;
                return CMD_BREAK;
            }
        }
    } else {
        if (scattering_distance <= s.distance) {
            if (use_weights) p.weight *= expf(-scattering_distance / s.absorption_length);
            p.time += scattering_distance / (CHROMA_SPEED_OF_LIGHT / s.n1);
            p.pos = p.pos + scattering_distance * p.dir;
            rayleigh_scatter(p, rng, K);
            p.history |= CHROMA_RAYLEIGH_SCATTER;
            p.last_slot = -1; p.last_tri = -1;
            return CMD_CONTINUE;
        }
    }

    if (use_weights) p.weight *= expf(-s.distance / s.absorption_length);
    p.pos = p.pos + s.distance * p.dir;
    p.time += s.distance / (CHROMA_SPEED_OF_LIGHT / s.n1);
    return CMD_PASS;
}

// photon.h:310-363
static __device__ __noinline__ void propagate_at_boundary(Photon &p, State &s, Rng &rng, const RngKey &K) {
    float incident_angle = get_theta(s.normal, -p.dir);
    float refracted_angle = asinf(sinf(incident_angle) * s.n1 / s.n2);

    float3 incident_plane_normal = cross3(p.dir, s.normal);
    float incident_plane_normal_length = norm3(incident_plane_normal);
    // at normal incidence the plane of incidence is not unique: use the polarisation
    if (incident_plane_normal_length < 1e-6f)
        incident_plane_normal = p.pol;
    else
        incident_plane_normal = incident_plane_normal / incident_plane_normal_length;

    float normal_coefficient = dot3(p.pol, incident_plane_normal);
    float normal_probability = normal_coefficient * normal_coefficient;

    float reflection_coefficient;
    if (rng_uniform(rng, K) < normal_probability) {
        // polarisation normal to the plane of incidence
        reflection_coefficient =
            -sinf(incident_angle - refracted_angle) / sinf(incident_angle + refracted_angle);
        if ((rng_uniform(rng, K) < reflection_coefficient * reflection_coefficient) ||
            isnan(refracted_angle)) {
            p.dir = rotate3(s.normal, incident_angle, incident_plane_normal);
            p.history |= CHROMA_REFLECT_SPECULAR;
        } else {
            p.dir = rotate3(s.normal, CHROMA_PI - refracted_angle, incident_plane_normal);
        }
        p.pol = incident_plane_normal;
    } else {
        // polarisation parallel to the plane of incidence
        reflection_coefficient =
            tanf(incident_angle - refracted_angle) / tanf(incident_angle + refracted_angle);
        if ((rng_uniform(rng, K) < reflection_coefficient * reflection_coefficient) ||
            isnan(refracted_angle)) {
            p.dir = rotate3(s.normal, incident_angle, incident_plane_normal);
            p.history |= CHROMA_REFLECT_SPECULAR;
        } else {
            p.dir = rotate3(s.normal, CHROMA_PI - refracted_angle, incident_plane_normal);
        }
        p.pol = cross3(incident_plane_normal, p.dir);
        p.pol = p.pol / norm3(p.pol);
    }
}

// photon.h:365-377
__device__ __forceinline__ int propagate_at_specular_reflector(Photon &p, State &s) {
    float incident_angle = get_theta(s.normal, -p.dir);
    float3 incident_plane_normal = cross3(p.dir, s.normal);
    incident_plane_normal = incident_plane_normal / norm3(incident_plane_normal);
    p.dir = rotate3(s.normal, incident_angle, incident_plane_normal);
    p.history |= CHROMA_REFLECT_SPECULAR;
    return CMD_CONTINUE;
}

// photon.h:379-398
__device__ __forceinline__ int propagate_at_diffuse_reflector(Photon &p, State &s, Rng &rng, const RngKey &K) {
    float ndotv;
    do {
        p.dir = uniform_sphere(rng, K);
        ndotv = dot3(p.dir, s.normal);
        if (ndotv < 0.0f) {
            p.dir = -p.dir;
            ndotv = -ndotv;
        }
    } while (!(rng_uniform(rng, K) < ndotv));
    p.pol = cross3(uniform_sphere(rng, K), p.dir);
    p.pol = p.pol / norm3(p.pol);
    p.history |= CHROMA_REFLECT_DIFFUSE;
    return CMD_CONTINUE;
}

// cx.h:24-34 helpers
// out of line: atan2f is ~60 instructions and the thin-film model takes eight of them (see philox4x32_10)
static __device__ __noinline__ float cx_arg2(float y, float x) { return atan2f(y, x); }
__device__ __forceinline__ float cx_arg(cuFloatComplex x) { return cx_arg2(x.y, x.x); }
__device__ __forceinline__ cuFloatComplex cx_sqrt(cuFloatComplex x) {
    float r = sqrtf(cuCabsf(x));
    float t = cx_arg(x) / 2.0f;
    return make_cuFloatComplex(r * cosf(t), r * sinf(t));
}

// one polarisation of the thin-film stack (photon.h:426-452 / 455-482 / 484-508):
// given r12, r23, t12, t23, g and the phase terms, return reflectance and transmittance
__device__ __forceinline__ void thin_film_rt(cuFloatComplex r12, cuFloatComplex r23,
                                             cuFloatComplex t12, cuFloatComplex t23,
                                             cuFloatComplex gg, float u, float v, float e,
                                             float &R, float &T) {
    float abs_r12 = cuCabsf(r12), abs_r23 = cuCabsf(r23);
    float abs_t12 = cuCabsf(t12), abs_t23 = cuCabsf(t23);
    float arg_r12 = cx_arg(r12), arg_r23 = cx_arg(r23);
    float exp1 = expf(2.0f * v * e);
    float exp2 = 1.0f / exp1;
    float denom = exp1 + abs_r12 * abs_r12 * abs_r23 * abs_r23 * exp2 +
                  2.0f * abs_r12 * abs_r23 * cosf(arg_r23 + arg_r12 + 2.0f * u * e);
    float r = abs_r12 * abs_r12 * exp1 + abs_r23 * abs_r23 * exp2 +
              2.0f * abs_r12 * abs_r23 * cosf(arg_r23 - arg_r12 + 2.0f * u * e);
    R = r / denom;
    float t = cuCrealf(gg) * abs_t12 * abs_t12 * abs_t23 * abs_t23;
    T = t / denom;
}

// photon.h:400-590
static __device__ __noinline__ int propagate_complex(Photon &p, State &s, Rng &rng, const RngKey &K, const GeomView &g,
                                              const float *s_opt, int si, bool use_weights) {
    const OpticsView &o = g.opt;
    const uint32_t base = o.surf_off + (uint32_t)si * 8u * o.nw;
    float detect = interp_property(o, p.wavelength, opt_tab(o, s_opt, base + 0 * o.nw));
    float reflect_diffuse = interp_property(o, p.wavelength, opt_tab(o, s_opt, base + 3 * o.nw));
    float n2_eta = interp_property(o, p.wavelength, opt_tab(o, s_opt, base + 5 * o.nw));
    float n2_k = interp_property(o, p.wavelength, opt_tab(o, s_opt, base + 6 * o.nw));
    const bool transmissive = o.surf_meta[si].y != 0;
    const float thickness = o.surf_thickness[si];

    // thin film optical model (RAT PMT optical model, P. Jones)
    cuFloatComplex n1 = make_cuFloatComplex(s.n1, 0.0f);
    cuFloatComplex n2 = make_cuFloatComplex(n2_eta, n2_k);
    cuFloatComplex n3 = make_cuFloatComplex(s.n2, 0.0f);
    const cuFloatComplex one = make_cuFloatComplex(1.0f, 0.0f);
    const cuFloatComplex two = make_cuFloatComplex(2.0f, 0.0f);

    float cos_t1 = dot3(p.dir, s.normal);
    if (cos_t1 < 0.0f) cos_t1 = -cos_t1;
    float theta = acosf(cos_t1);
    cuFloatComplex cos1 = make_cuFloatComplex(cosf(theta), 0.0f);
    cuFloatComplex sin1 = make_cuFloatComplex(sinf(theta), 0.0f);

    float e = 2.0f * CHROMA_PI * thickness * 1.0e6 / p.wavelength;  // thickness mm -> nm
    cuFloatComplex ratio13sin = cuCmulf(cuCmulf(cuCdivf(n1, n3), cuCdivf(n1, n3)), cuCmulf(sin1, sin1));
    cuFloatComplex cos3 = cx_sqrt(cuCsubf(one, ratio13sin));
    cuFloatComplex ratio12sin = cuCmulf(cuCmulf(cuCdivf(n1, n2), cuCdivf(n1, n2)), cuCmulf(sin1, sin1));
    cuFloatComplex cos2 = cx_sqrt(cuCsubf(one, ratio12sin));
    float u = cuCrealf(cuCmulf(n2, cos2));
    float v = cuCimagf(cuCmulf(n2, cos2));

    // s polarisation
    float s_r, s_t;
    {
        cuFloatComplex n1c1 = cuCmulf(n1, cos1), n2c2 = cuCmulf(n2, cos2), n3c3 = cuCmulf(n3, cos3);
        cuFloatComplex r12 = cuCdivf(cuCsubf(n1c1, n2c2), cuCaddf(n1c1, n2c2));
        cuFloatComplex r23 = cuCdivf(cuCsubf(n2c2, n3c3), cuCaddf(n2c2, n3c3));
        cuFloatComplex t12 = cuCdivf(cuCmulf(two, n1c1), cuCaddf(n1c1, n2c2));
        cuFloatComplex t23 = cuCdivf(cuCmulf(two, n2c2), cuCaddf(n2c2, n3c3));
        cuFloatComplex gg = cuCdivf(n3c3, n1c1);
        thin_film_rt(r12, r23, t12, t23, gg, u, v, e, s_r, s_t);
    }
    // p polarisation
    float p_r, p_t;
    {
        cuFloatComplex n2c1 = cuCmulf(n2, cos1), n3c2 = cuCmulf(n3, cos2);
        cuFloatComplex n2c3 = cuCmulf(n2, cos3), n1c2 = cuCmulf(n1, cos2);
        cuFloatComplex r12 = cuCdivf(cuCsubf(n2c1, n1c2), cuCaddf(n2c1, n1c2));
        cuFloatComplex r23 = cuCdivf(cuCsubf(n3c2, n2c3), cuCaddf(n3c2, n2c3));
        cuFloatComplex t12 = cuCdivf(cuCmulf(cuCmulf(two, n1), cos1), cuCaddf(n2c1, n1c2));
        cuFloatComplex t23 = cuCdivf(cuCmulf(cuCmulf(two, n2), cos2), cuCaddf(n3c2, n2c3));
        cuFloatComplex gg = cuCdivf(cuCmulf(n3, cos3), cuCmulf(n1, cos1));
        thin_film_rt(r12, r23, t12, t23, gg, u, v, e, p_r, p_t);
    }
    // normal incidence (for the QE rescale)
    float n_r, n_t;
    {
        cuFloatComplex r12 = cuCdivf(cuCsubf(n1, n2), cuCaddf(n1, n2));
        cuFloatComplex r23 = cuCdivf(cuCsubf(n2, n3), cuCaddf(n2, n3));
        cuFloatComplex t12 = cuCdivf(cuCmulf(two, n1), cuCaddf(n1, n2));
        cuFloatComplex t23 = cuCdivf(cuCmulf(two, n2), cuCaddf(n2, n3));
        cuFloatComplex gg = cuCdivf(n3, n1);
        thin_film_rt(r12, r23, t12, t23, gg, n2_eta, n2_k, e, n_r, n_t);
    }

    // s polarisation fraction, as in propagate_at_boundary
    float incident_angle = get_theta(s.normal, -p.dir);
    float refracted_angle = asinf(sinf(incident_angle) * s.n1 / s.n2);
    float3 incident_plane_normal = cross3(p.dir, s.normal);
    float incident_plane_normal_length = norm3(incident_plane_normal);
    if (incident_plane_normal_length < 1e-6f)
        incident_plane_normal = p.pol;
    else
        incident_plane_normal = incident_plane_normal / incident_plane_normal_length;
    float normal_coefficient = dot3(p.pol, incident_plane_normal);
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

    detect /= absorb_normal_incidence;  // scale detection efficiency on absorption
    if (use_weights && p.weight > WEIGHT_LOWER_THRESHOLD && absorb < (1.0f - WEIGHT_LOWER_THRESHOLD)) {
        float survive = 1.0f - absorb;
        absorb = 0.0f;
        p.weight *= survive;
        detect /= survive;
        reflect /= survive;
        transmit /= survive;
    }
    if (use_weights && detect > 0.0f) {
        p.history |= CHROMA_SURFACE_DETECT;
        p.weight *= detect;
        return CMD_BREAK;
    }

    float uniform_sample = rng_uniform(rng, K);
    if (uniform_sample < absorb) {
        // detection probability is conditional on absorption here
        float uniform_sample_detect = rng_uniform(rng, K);
        if (uniform_sample_detect < detect) p.history |= CHROMA_SURFACE_DETECT;
        else p.history |= CHROMA_SURFACE_ABSORB;
        return CMD_BREAK;
    } else if (uniform_sample < absorb + reflect || !transmissive) {
        float uniform_sample_reflect = rng_uniform(rng, K);
        if (uniform_sample_reflect < reflect_diffuse) return propagate_at_diffuse_reflector(p, s, rng, K);
        else return propagate_at_specular_reflector(p, s);
    } else {
        p.dir = rotate3(s.normal, CHROMA_PI - refracted_angle, incident_plane_normal);
        p.pol = cross3(incident_plane_normal, p.dir);
        p.pol = p.pol / norm3(p.pol);
        p.history |= CHROMA_SURFACE_TRANSMIT;
        return CMD_CONTINUE;
    }
}

// photon.h:592-637
__device__ __forceinline__ int propagate_at_wls(Photon &p, State &s, Rng &rng, const RngKey &K, const GeomView &g,
                                                const float *s_opt, int si, bool use_weights) {
    const OpticsView &o = g.opt;
    const uint32_t base = o.surf_off + (uint32_t)si * 8u * o.nw;
    float absorb = interp_property(o, p.wavelength, opt_tab(o, s_opt, base + 1 * o.nw));
    float reflect_specular = interp_property(o, p.wavelength, opt_tab(o, s_opt, base + 4 * o.nw));
    float reflect_diffuse = interp_property(o, p.wavelength, opt_tab(o, s_opt, base + 3 * o.nw));
    float reemit = interp_property(o, p.wavelength, opt_tab(o, s_opt, base + 2 * o.nw));

    float uniform_sample = rng_uniform(rng, K);

    if (use_weights && p.weight > WEIGHT_LOWER_THRESHOLD && absorb < (1.0f - WEIGHT_LOWER_THRESHOLD)) {
        float survive = 1.0f - absorb;
        absorb = 0.0f;
        p.weight *= survive;
        reflect_diffuse /= survive;
        reflect_specular /= survive;
    }

    if (uniform_sample < absorb) {
        float uniform_sample_reemit = rng_uniform(rng, K);
        if (uniform_sample_reemit < reemit) {
            p.history |= CHROMA_SURFACE_REEMIT;
            p.wavelength = sample_cdf_uniform(rng, K, o.nw, o.w0, o.dw, opt_tab(o, s_opt, base + 7 * o.nw));
            p.dir = uniform_sphere(rng, K);
            p.pol = cross3(uniform_sphere(rng, K), p.dir);
            p.pol = p.pol / norm3(p.pol);
            return CMD_CONTINUE;
        } else {
            p.history |= CHROMA_SURFACE_ABSORB;
            return CMD_BREAK;
        }
    } else if (uniform_sample < absorb + reflect_specular + reflect_diffuse) {
        float uniform_sample_reflect = rng_uniform(rng, K) * (reflect_specular + reflect_diffuse);
        if (uniform_sample_reflect < reflect_specular) return propagate_at_specular_reflector(p, s);
        else return propagate_at_diffuse_reflector(p, s, rng, K);
    } else {
        p.history |= CHROMA_SURFACE_TRANSMIT;
        return CMD_PASS;
    }
}

// photon.h:640-670
static __device__ __noinline__ int propagate_at_dichroic(Photon &p, State &s, Rng &rng, const RngKey &K, const GeomView &g,
                                                  int si) {
    const OpticsView &o = g.opt;
    float incident_angle = get_theta(s.normal, -p.dir);
    uint4 meta = o.surf_meta[si];
    const uint32_t nangles = meta.w;
    const float *angles = o.blob + meta.z;
    const float *refl = angles + nangles;
    const float *tran = refl + (size_t)nangles * o.nw;
    float idx = interp_idx(incident_angle, nangles, angles);
    unsigned int iidx = (int)idx;
    unsigned int iidx_hi = iidx < nangles - 2 ? iidx + 1 : iidx;
    float reflect_prob_low = interp_property(o, p.wavelength, refl + (size_t)iidx * o.nw);
    float reflect_prob_high = interp_property(o, p.wavelength, refl + (size_t)iidx_hi * o.nw);
    float transmit_prob_low = interp_property(o, p.wavelength, tran + (size_t)iidx * o.nw);
    float transmit_prob_high = interp_property(o, p.wavelength, tran + (size_t)iidx_hi * o.nw);
    float reflect_prob = reflect_prob_low + (reflect_prob_high - reflect_prob_low) * (idx - iidx);
    float transmit_prob = transmit_prob_low + (transmit_prob_high - transmit_prob_low) * (idx - iidx);

    float uniform_sample = rng_uniform(rng, K);
    if (uniform_sample < reflect_prob) {
        return propagate_at_specular_reflector(p, s);
    } else if (uniform_sample < transmit_prob + reflect_prob) {
        p.history |= CHROMA_SURFACE_TRANSMIT;
        return CMD_PASS;
    } else {
        p.history |= CHROMA_SURFACE_ABSORB;
        return CMD_BREAK;
    }
}

// photon.h:672-733
__device__ __forceinline__ int propagate_at_surface(Photon &p, State &s, Rng &rng, const RngKey &K,
                                                    const GeomView &g, const float *s_opt,
                                                    bool use_weights) {
    const OpticsView &o = g.opt;
    const int si = s.surface_index;
    const uint32_t model = o.surf_meta[si].x;
    if (model == CHROMA_SURFACE_COMPLEX) return propagate_complex(p, s, rng, K, g, s_opt, si, use_weights);
    if (model == CHROMA_SURFACE_WLS) return propagate_at_wls(p, s, rng, K, g, s_opt, si, use_weights);
    if (model == CHROMA_SURFACE_DICHROIC) return propagate_at_dichroic(p, s, rng, K, g, si);

    // default model: the four probabilities are linearly interpolated so they still sum to 1
    const uint32_t base = o.surf_off + (uint32_t)si * 8u * o.nw;
    float detect = interp_property(o, p.wavelength, opt_tab(o, s_opt, base + 0 * o.nw));
    float absorb = interp_property(o, p.wavelength, opt_tab(o, s_opt, base + 1 * o.nw));
    float reflect_diffuse = interp_property(o, p.wavelength, opt_tab(o, s_opt, base + 3 * o.nw));
    float reflect_specular = interp_property(o, p.wavelength, opt_tab(o, s_opt, base + 4 * o.nw));

    float uniform_sample = rng_uniform(rng, K);

    if (use_weights && p.weight > WEIGHT_LOWER_THRESHOLD && absorb < (1.0f - WEIGHT_LOWER_THRESHOLD)) {
        float survive = 1.0f - absorb;
        absorb = 0.0f;
        p.weight *= survive;
        detect /= survive;
        reflect_diffuse /= survive;
        reflect_specular /= survive;
    }
    if (use_weights && detect > 0.0f) {
        p.history |= CHROMA_SURFACE_DETECT;
        p.weight *= detect;
        return CMD_BREAK;
    }
    if (uniform_sample < absorb) {
        p.history |= CHROMA_SURFACE_ABSORB;
        return CMD_BREAK;
    } else if (uniform_sample < absorb + detect) {
        p.history |= CHROMA_SURFACE_DETECT;
        return CMD_BREAK;
    } else if (uniform_sample < absorb + detect + reflect_diffuse) {
        return propagate_at_diffuse_reflector(p, s, rng, K);
    } else if (uniform_sample < absorb + detect + reflect_diffuse + reflect_specular) {
        return propagate_at_specular_reflector(p, s);
    } else {
        return CMD_PASS;
    }
}
