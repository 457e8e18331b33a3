/* oracle/ref_cpu_wrapper.cpp — TEST INFRASTRUCTURE ONLY (checker + timed CPU baseline).
 *
 * Compiles the reference's OWN sources for the host, from where they lie under /root/reference
 * (nothing is copied): chroma/cuda/propagate.cu (with photon.h, mesh.h, intersect.h, geometry.h,
 * random.h, rotate.h, cx.h ...) and chroma/cuda/daq.cu are #included below as plain C++.
 * What makes that possible:
 *   - CUDA's vector_types.h / vector_functions.h / cuComplex.h are host-compilable headers;
 *   - the execution-space qualifiers are defined away, threadIdx/blockIdx/blockDim become
 *     thread-local variables, __syncthreads() a no-op, the atomics GCC builtins;
 *   - <curand_kernel.h> resolves to oracle/ref_cpu/curand_kernel.h: a Philox stream with the
 *     engine's keying, so the reference code draws the same uniforms as the engine / oracle port.
 * A "thread" is run by setting blockIdx.x = photon index, blockDim.x = 1, threadIdx.x = 0 and
 * calling the reference's __global__ function itself (propagate, run_daq, ...), so the kernel
 * bodies execute unmodified.  Arithmetic is IEEE (no --use_fast_math, -ffp-contract=off): this is
 * the reference source's exact-arithmetic reading, not its sm_100a cubin (that one runs on the
 * GPU box through oracle/ref_gpu.py).
 *
 * Built by oracle/Makefile (target ref_cpu) into oracle/_ref/libchroma_ref_cpu.so. */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <cmath>
#include <vector>
#include <omp.h>

#include <vector_types.h>
#include <vector_functions.h>
#include <cuComplex.h>

#undef __device__
#undef __global__
#undef __host__
#undef __shared__
#undef __noinline__
#define __device__
#define __global__
#define __host__
#define __shared__
#define __noinline__

using std::isnan;
using std::isfinite;

static thread_local uint3 threadIdx = {0, 0, 0}, blockIdx = {0, 0, 0};
static thread_local uint3 blockDim = {1, 1, 1}, gridDim = {1, 1, 1};
static inline void __syncthreads() {}
static inline float __int_as_float(int i) { float f; memcpy(&f, &i, 4); return f; }
static inline int __float_as_int(float f) { int i; memcpy(&i, &f, 4); return i; }
static inline float max(float a, float b) { return fmaxf(a, b); }
static inline float min(float a, float b) { return fminf(a, b); }
static inline unsigned int atomicAdd(unsigned int *p, unsigned int v) { return __atomic_fetch_add(p, v, __ATOMIC_RELAXED); }
static inline int atomicAdd(int *p, int v) { return __atomic_fetch_add(p, v, __ATOMIC_RELAXED); }
static inline unsigned int atomicOr(unsigned int *p, unsigned int v) { return __atomic_fetch_or(p, v, __ATOMIC_RELAXED); }
static inline unsigned int atomicMin(unsigned int *p, unsigned int v) {
    unsigned int old = __atomic_load_n(p, __ATOMIC_RELAXED);
    while (v < old && !__atomic_compare_exchange_n(p, &old, v, true, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {}
    return old;
}

// ---- the reference's own sources ---------------------------------------------------------
#include "propagate.cu"   // chroma/cuda/propagate.cu
#include "daq.cu"         // chroma/cuda/daq.cu
// -------------------------------------------------------------------------------------------

#include "chroma_oracle.h"   // oracle_geometry: the flat description both checkers take

namespace {

// The reference's pointer-linked device structs (geometry_types.h), assembled from the flat tables
// the way chroma/gpu/geometry.py:13-259 assembles them on the device.
struct RefGeometry {
    Geometry g;
    std::vector<Material> materials;
    std::vector<Material *> material_ptrs;
    std::vector<std::vector<float *>> comp_ptrs;   // per material: 4 lists of per-component pointers
    std::vector<Surface> surfaces;
    std::vector<Surface *> surface_ptrs;
    std::vector<DichroicProps> dichroic;
    std::vector<std::vector<float *>> dich_ptrs;
    Detector det;

    explicit RefGeometry(const oracle_geometry *o) {
        const uint32_t nw = o->nwavelengths;
        memset(&g, 0, sizeof(g));
        g.vertices = (float3 *)o->vertices;
        g.triangles = (uint3 *)o->triangles;
        g.material_codes = (unsigned int *)o->material_codes;
        g.primary_nodes = (uint4 *)o->nodes;
        g.extra_nodes = nullptr;
        g.nprimary_nodes = (int)o->nnodes;
        g.world_origin = make_float3(o->world_origin[0], o->world_origin[1], o->world_origin[2]);
        g.world_scale = o->world_scale;
        materials.resize(o->nmaterials);
        comp_ptrs.resize((size_t)o->nmaterials * 4);
        uint32_t c0 = 0;
        for (uint32_t m = 0; m < o->nmaterials; m++) {
            Material &M = materials[m];
            float *base = (float *)o->material_props + (size_t)m * 3 * nw;
            M.refractive_index = base; M.absorption_length = base + nw; M.scattering_length = base + 2 * nw;
            const uint32_t nc = o->material_num_comp ? o->material_num_comp[m] : 0;
            for (uint32_t c = 0; c < nc; c++) {
                comp_ptrs[m * 4 + 0].push_back((float *)o->comp_reemission_prob + (size_t)(c0 + c) * nw);
                comp_ptrs[m * 4 + 1].push_back((float *)o->comp_reemission_wvl_cdf + (size_t)(c0 + c) * nw);
                comp_ptrs[m * 4 + 2].push_back((float *)o->comp_reemission_time_cdf + (size_t)(c0 + c) * o->ntimes);
                comp_ptrs[m * 4 + 3].push_back((float *)o->comp_absorption_length + (size_t)(c0 + c) * nw);
            }
            c0 += nc;
            M.comp_reemission_prob = comp_ptrs[m * 4 + 0].data();
            M.comp_reemission_wvl_cdf = comp_ptrs[m * 4 + 1].data();
            M.comp_reemission_time_cdf = comp_ptrs[m * 4 + 2].data();
            M.comp_absorption_length = comp_ptrs[m * 4 + 3].data();
            M.num_comp = nc;
            M.wavelength_n = nw; M.wavelength_step = o->wavelength_step; M.wavelength_start = o->wavelength_start;
            M.time_n = o->ntimes; M.time_step = o->time_step; M.time_start = o->time_start;
        }
        for (auto &M : materials) material_ptrs.push_back(&M);
        g.materials = material_ptrs.data();
        surfaces.resize(o->nsurfaces);
        dichroic.resize(o->nsurfaces);
        dich_ptrs.resize((size_t)o->nsurfaces * 2);
        uint64_t a0 = 0;
        for (uint32_t s = 0; s < o->nsurfaces; s++) {
            Surface &S = surfaces[s];
            memset(&S, 0, sizeof(S));
            float *base = (float *)o->surface_props + (size_t)s * 8 * nw;
            S.detect = base; S.absorb = base + nw; S.reemit = base + 2 * nw; S.reflect_diffuse = base + 3 * nw;
            S.reflect_specular = base + 4 * nw; S.eta = base + 5 * nw; S.k = base + 6 * nw;
            S.reemission_cdf = base + 7 * nw;
            S.model = o->surface_model[s];
            S.transmissive = o->surface_transmissive[s];
            S.thickness = o->surface_thickness[s];
            S.wavelength_n = nw; S.wavelength_step = o->wavelength_step; S.wavelength_start = o->wavelength_start;
            const uint32_t na = o->dichroic_nangles ? o->dichroic_nangles[s] : 0;
            if (na) {
                DichroicProps &D = dichroic[s];
                D.angles = (float *)o->dichroic_angles + a0;
                for (uint32_t a = 0; a < na; a++) {
                    dich_ptrs[s * 2 + 0].push_back((float *)o->dichroic_reflect + (a0 + a) * nw);
                    dich_ptrs[s * 2 + 1].push_back((float *)o->dichroic_transmit + (a0 + a) * nw);
                }
                D.dichroic_reflect = dich_ptrs[s * 2 + 0].data();
                D.dichroic_transmit = dich_ptrs[s * 2 + 1].data();
                D.nangles = na;
                S.dichroic_props = &D;
            }
            a0 += na;
        }
        for (uint32_t s = 0; s < o->nsurfaces; s++)
            surface_ptrs.push_back((!o->surface_present || o->surface_present[s]) ? &surfaces[s] : nullptr);
        g.surfaces = surface_ptrs.data();
        memset(&det, 0, sizeof(det));
        if (o->nchannels) {
            det.solid_id_to_channel_index = (int *)o->solid_id_to_channel_index;
            det.time_cdf_x = (float *)o->time_cdf_x; det.time_cdf_y = (float *)o->time_cdf_y;
            det.charge_cdf_x = (float *)o->charge_cdf_x; det.charge_cdf_y = (float *)o->charge_cdf_y;
            det.nchannels = (int)o->nchannels;
            det.time_cdf_len = (int)o->time_cdf_len; det.charge_cdf_len = (int)o->charge_cdf_len;
            det.charge_unit = o->charge_cdf_x[o->charge_cdf_len - 1] / 65536.0f;   // chroma/gpu/detector.py:39
        }
    }
};

inline void as_thread(uint64_t id) {
    blockIdx.x = (unsigned)id; blockDim.x = 1; threadIdx.x = 0;
}

}  // namespace

extern "C" int ref_cpu_num_threads(void) { return omp_get_max_threads(); }

/* chroma/cuda/mesh.h:41-151: intersect_mesh per ray, direction normalised as distance_to_mesh does */
extern "C" void ref_cpu_intersect(const oracle_geometry *o, uint64_t n, const float *pos, const float *dir,
                                  const int32_t *last_hit, int32_t *out_tri, float *out_dist) {
    RefGeometry R(o);
#pragma omp parallel for schedule(dynamic, 256)
    for (int64_t id = 0; id < (int64_t)n; id++) {
        float3 origin = make_float3(pos[3 * id], pos[3 * id + 1], pos[3 * id + 2]);
        float3 direction = make_float3(dir[3 * id], dir[3 * id + 1], dir[3 * id + 2]);
        direction /= norm(direction);
        float distance;
        int tri = intersect_mesh(origin, direction, &R.g, distance, last_hit ? last_hit[id] : -1);
        out_tri[id] = tri;
        out_dist[id] = tri != -1 ? distance : -1.0f;
    }
}

/* chroma/cuda/propagate.cu:217-319: the reference's `propagate` kernel, one call per photon */
extern "C" void ref_cpu_propagate(const oracle_geometry *o, uint64_t n, float *pos, float *dir, float *pol,
                                  float *wavelengths, float *t, int32_t *last_hit, uint32_t *flags,
                                  float *weights, uint32_t *evidx, uint64_t seed, uint32_t id_base,
                                  uint32_t epoch, int max_steps, int use_weights, int scatter_first) {
    RefGeometry R(o);
    std::vector<curandState> states(n);
    std::vector<unsigned int> in_queue(n), out_queue(n + 1);
    for (uint64_t i = 0; i < n; i++) {
        refcpu_rng_seed(&states[i], seed, id_base + (uint32_t)i, epoch, 0u);
        in_queue[i] = (unsigned)i;
    }
    out_queue[0] = 1;   // chroma/gpu/photon.py:243
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t id = 0; id < (int64_t)n; id++) {
        as_thread((uint64_t)id);
        propagate(0, (int)n, in_queue.data(), out_queue.data(), states.data(), (float3 *)pos, (float3 *)dir,
                  wavelengths, (float3 *)pol, t, flags, last_hit, weights, evidx, max_steps, use_weights,
                  scatter_first, &R.g);
    }
}

/* chroma/cuda/daq.cu:25-172: reset_earliest_time_int, run_daq, convert_* — the sequence of
 * chroma/gpu/daq.py:54-99 (begin_acquire / acquire / end_acquire) */
extern "C" void ref_cpu_run_daq(const oracle_geometry *o, uint64_t n, float *t, uint32_t *flags, int32_t *last_hit,
                                float *weights, uint64_t seed, uint32_t id_base, uint32_t epoch,
                                float global_weight, float *out_t, float *out_q, uint32_t *out_hist) {
    RefGeometry R(o);
    const int nch = R.det.nchannels;
    std::vector<unsigned int> time_int(nch), q_int(nch, 0u);
    for (int i = 0; i < nch; i++) { as_thread(i); reset_earliest_time_int(1e9f, nch, time_int.data()); }
    for (int i = 0; i < nch; i++) out_hist[i] = 0u;
    std::vector<curandState> states(n);
    for (uint64_t i = 0; i < n; i++) refcpu_rng_seed(&states[i], seed, id_base + (uint32_t)i, epoch, 1u);
    for (uint64_t id = 0; id < n; id++) {
        as_thread(id);
        run_daq(states.data(), 0x1 << 2, 0, (int)n, t, flags, last_hit, weights, (int *)o->solid_id, &R.det,
                time_int.data(), q_int.data(), out_hist, global_weight);
    }
    for (int i = 0; i < nch; i++) {
        as_thread(i);
        convert_sortable_int_to_float(nch, time_int.data(), out_t);
        convert_charge_int_to_float(&R.det, q_int.data(), out_q);
    }
}
