// common.cuh — shared host/device declarations of libchroma_b200 (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/chroma_b200.h"

// ---------------------------------------------------------------------------------------
// error handling
// ---------------------------------------------------------------------------------------
void chroma_set_error(const char *fmt, ...);

#define CU_CHECK(expr)                                                                     \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess) {                                                           \
            chroma_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,                 \
                             cudaGetErrorString(_e));                                      \
            return CHROMA_ERR_CUDA;                                                        \
        }                                                                                  \
    } while (0)

#define ARG_CHECK(cond)                                                                    \
    do {                                                                                   \
        if (!(cond)) {                                                                     \
            chroma_set_error("%s:%d: bad argument: %s", __FILE__, __LINE__, #cond);        \
            return CHROMA_ERR_ARG;                                                         \
        }                                                                                  \
    } while (0)

// ---------------------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------------------
struct chroma_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    int sm_count = 0;
    size_t l2_bytes = 0, l2_persist_max = 0;
    int cc = 0;
    uint64_t launches = 0;
    void *flush_buf = nullptr;
    size_t flush_bytes = 0;
    // small device scratch (counters / status words), zeroed on demand
    uint32_t *d_scratch = nullptr;  // 64 words
    uint32_t *h_scratch = nullptr;  // pinned mirror
    // side streams of the wavefront propagate (engine_wavefront.cu), created on first use
    cudaStream_t wave_stream[16] = {};
    cudaEvent_t wave_event[16] = {};
    cudaEvent_t wave_fork = nullptr;
    int n_wave_streams = 0;
    // copy stream of the chunked photon upload (engine_photons.cu), created with the context
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t copy_fork = nullptr;
    // page-locked staging ring for uploads from PAGEABLE host arrays (engine_photons.cu): slot k holds one
    // chunk of every field; a slot is reused once its DMA has completed (slot_done[k])
    char *h_stage_ring = nullptr;
    size_t stage_slot_bytes = 0;
    cudaEvent_t stage_slot_done[4] = {nullptr, nullptr, nullptr, nullptr};
    uint32_t stage_slot_pos = 0;
    // page-locked ring of progress values, the sources of the 8-byte "photons landed" DMA writes
    unsigned long long *h_ready_ring = nullptr;
    uint32_t ready_ring_pos = 0;
    // worker thread (std::thread *) of an upload whose pageable sources were lent to the engine
    // (chroma_photon_arrays.borrow_sources), joined by uploader_join()
    void *uploader = nullptr;
    int uploader_rc = 0;
    char uploader_err[512] = "";
    // per-lane photon slots of the two-slot propagate kernel (wide_prop2.cuh), allocated on first use
    uint32_t *d_side = nullptr;
    size_t side_bytes = 0;
    // tuning knobs, read from the environment ONCE when the context is created
    // (chroma_ctx_reload_knobs re-reads them: tools/prop_sweep.py)
    struct Knobs {
        int refill = 20, iters = 16, leaf_thresh = 8;       // propagate: CHROMA_REFILL, CHROMA_WIDE_ITERS, CHROMA_WIDE_LEAF_THRESH
        int isect_refill = 1, isect_iters = 8;              // intersect: CHROMA_ISECT_REFILL, CHROMA_ISECT_ITERS
        int l2_window_mb = 64;                              // CHROMA_L2_WINDOW_MB (0 = off)
        int upload_direct = 1;                              // CHROMA_UPLOAD_DIRECT
        int prop_v = 4;                                     // CHROMA_PROP_V (A/B kernels)
        int isect_v = 3;                                    // CHROMA_ISECT_V: 1 = reference-order search (debug)
        int count_wide = 0;                                 // CHROMA_COUNT_WIDE: counting entry on the engine's tree
        int upload_chunks = 16;                             // CHROMA_UPLOAD_CHUNKS (0 = plain synchronous copy)
        int upload_threads = 8;                             // CHROMA_UPLOAD_THREADS: host threads that fill the pinned ring
        int strict = 0;                                     // CHROMA_STRICT: no silent slow paths
        int rng_ool = -1;                                   // CHROMA_RNG_OOL: -1 by geometry, 0 | 1 force the kernel build
        unsigned watchdog_spins = 4000000u;                 // CHROMA_UPLOAD_WATCHDOG: ~us the kernel waits for staged photons
    } knobs;
    // per-device, per-context function attributes already set (dynamic shared memory opt-in)
    bool wide_attr_set = false, inst_attr_set = false, wide_ool_attr_set = false, inst_ool_attr_set = false;
    // sticky status of asynchronous failures (search-stack overflow, stalled upload): reported by the
    // next synchronising call even when the launching call was not asked for its status
    uint32_t sticky_status = 0;
    bool status_pending = false;       // a propagate launch whose status word has not been read yet
    struct chroma_geometry *l2_window_geom = nullptr;   // geometry the stream's access-policy window points into
};
int chroma_status_to_error(uint32_t status);
// A/B kernels of round 1 (reference-order searches v1-v3, two photons per lane): compiled only into
// variant builds (python -m chroma_b200.csrc.build --variant ab -DCHROMA_AB_KERNELS=1)
#ifndef CHROMA_AB_KERNELS
#define CHROMA_AB_KERNELS 0
#endif
void chroma_ctx_read_knobs(chroma_ctx *c);
#define READY_RING 4096

static inline int ctx_bind(chroma_ctx *ctx) {
    if (!ctx) {
        chroma_set_error("null context");
        return CHROMA_ERR_ARG;
    }
    CU_CHECK(cudaSetDevice(ctx->device));
    return 0;
}

#define CTX_BIND(ctx)                                                                      \
    do {                                                                                   \
        int _r = ctx_bind(ctx);                                                            \
        if (_r) return _r;                                                                 \
    } while (0)

#define KERNEL_CHECK(ctx)                                                                  \
    do {                                                                                   \
        (ctx)->launches++;                                                                 \
        CU_CHECK(cudaGetLastError());                                                      \
    } while (0)

static inline unsigned grid_for(uint64_t n, unsigned block, chroma_ctx *ctx, unsigned per_sm) {
    uint64_t need = (n + block - 1) / block;
    uint64_t cap = (uint64_t)ctx->sm_count * per_sm;
    if (need < 1) need = 1;
    return (unsigned)(need < cap ? need : cap);
}

// ---------------------------------------------------------------------------------------
// device-side geometry view (passed to kernels by value)
// ---------------------------------------------------------------------------------------
// Optics tables live in one float blob.  Offsets are in floats from the blob start.
//   [mat_off  + (m*3 + k)*nw]            k: 0 refractive_index 1 absorption 2 scattering
//   [comp_off + (c*3 + k)*nw]            k: 0 reemission_prob 1 reemission_wvl_cdf 2 comp_abs
//   [surf_off + (s*8 + k)*nw]            k: detect absorb reemit rdiffuse rspecular eta k rcdf
//   [dich_off[s] ...]                    angles[na], reflect[na][nw], transmit[na][nw]
//   [tcdf_off + c*nt]                    reemission time CDFs (never staged in smem)
struct OpticsView {
    const float *blob;        // global memory
    uint32_t smem_floats;     // leading part of the blob staged in shared memory (0 = none)
    uint32_t nw, nt;
    float w0, dw, t0, dt;
    uint32_t mat_off, comp_off, surf_off, tcdf_off;
    uint32_t nmat, nsurf;
    const uint32_t *mat_ncomp;      // [nmat]
    const uint32_t *mat_comp_first; // [nmat]
    // per surface: x=model y=transmissive z=dichroic offset in blob (0xFFFFFFFF none) w=nangles
    const uint4 *surf_meta;         // [nsurf]
    const float *surf_thickness;    // [nsurf]
};

struct GeomView {
    const uint4 *nodes;      // engine node array: reference packing, leaf w = slot
    const float4 *tris;      // 3 x float4 per leaf slot (see pack_triangles)
    const int32_t *tri_slot; // triangle id -> slot (-1 unreachable)
    float3 world_origin;
    float world_scale;
    uint32_t nnodes, nslots, ntriangles;
    uint32_t top_nodes;      // first `top_nodes` nodes are staged in shared memory
    OpticsView opt;
    // instanced geometry (engine_inst.cu): no flat arrays; a triangle's solid is found in tri_base
    const uint32_t *inst_tri_base;   // [ninst + 1] or NULL
    uint32_t ninst;
};

// the engine's own search structure (wide.cuh, engine_wide.cu)
struct WideView {
    const uint4 *groups;     // [ngroups][8] records, one 128-byte line per group
    const float4 *tris;      // [nslots][3] packed triangle records, Morton order
    const uint32_t *rank;    // [nslots] reference search rank (tie break)
    uint32_t ngroups, nslots;
    float3 world_origin;
    float world_scale;
};

// two-level tree of placed templates (inst.cuh, engine_inst.cu)
struct InstView {
    const uint4 *groups;
    const float4 *tris;
    const float4 *inst;
    const uint32_t *tri_base;          // [ninst + 1] first global triangle id per solid (add order)
    const uint32_t *tmpl_off_of_orig;  // [ninst] solid id -> offset of its template in tris / tri_slot
    const int32_t *tri_slot;           // [template triangles] offset + triangle id inside the solid -> slot
    uint32_t ngroups, ninst, ntriangles, nslots;
    uint32_t top_root;                 // first group of the top tree (the template trees come first)
    float3 top_origin;
    float top_scale;
    float guard_lim;                   // |origin| beyond this cannot be searched (wide.cuh wide_origin_unsearchable)
};

struct DetectorView {
    const int32_t *solid_to_channel;
    const float *time_cdf_x, *time_cdf_y, *charge_cdf_x, *charge_cdf_y;
    int nchannels, time_cdf_len, charge_cdf_len;
    float charge_unit;
};

// photon SoA-of-float4 ("AoSoA") device state: four 16-byte records per photon
struct PhotonView {
    float4 *pt;   // pos.xyz, t
    float4 *dw;   // dir.xyz, wavelength
    float4 *pw;   // pol.xyz, weight
    uint4 *st;    // flags, last_hit_triangle (as uint), evidx, channel (hits only)
};

#define TERMINAL_MASK                                                                      \
    (CHROMA_NO_HIT | CHROMA_BULK_ABSORB | CHROMA_SURFACE_DETECT | CHROMA_SURFACE_ABSORB |  \
     (unsigned)CHROMA_NAN_ABORT)

// host-side objects --------------------------------------------------------------------
struct chroma_bvh {
    chroma_ctx *ctx = nullptr;
    uint4 *d_nodes = nullptr;
    uint64_t nnodes = 0;
    std::vector<uint64_t> layer_offsets;
    float world_origin[3] = {0, 0, 0};
    float world_scale = 1.f;
};

struct chroma_geometry {
    chroma_ctx *ctx = nullptr;
    // reference-format arrays kept for API parity (GPUGeometry attributes)
    uint4 *d_nodes_ref = nullptr;
    float *d_vertices = nullptr;
    uint32_t *d_triangles = nullptr;
    uint32_t *d_material_codes = nullptr;
    uint32_t *d_colors = nullptr;
    uint32_t *d_solid_id = nullptr;
    uint32_t *d_leaf_bits = nullptr;   // 1 bit per node: reachable leaf (lets d_nodes_ref be rebuilt from d_nodes)
    bool mesh_released = false;        // chroma_geometry_release_mesh: vertices / triangles / codes / solid ids freed
    // engine-internal arrays
    uint4 *d_nodes = nullptr;
    float4 *d_tris = nullptr;
    int32_t *d_tri_slot = nullptr;
    float *d_optics = nullptr;
    uint32_t *d_mat_ncomp = nullptr, *d_mat_comp_first = nullptr;
    uint4 *d_surf_meta = nullptr;
    float *d_surf_thickness = nullptr;
    uint64_t nnodes = 0, ntriangles = 0, nvertices = 0, nslots = 0;
    uint64_t bytes = 0;
    GeomView view;
    // wide tree: arrays in Morton slot order; wview = view with tris / tri_slot of that order
    uint4 *d_wgroups = nullptr;
    float4 *d_wtris = nullptr;
    uint32_t *d_wrank = nullptr;
    int32_t *d_wtri_slot = nullptr;
    WideView wide{};
    GeomView wview;
    uint32_t wide_levels = 0;
    uint32_t wide_leaves = 0;
    // reemitting materials or thin-film / WLS / dichroic surfaces: the physics, not the search, dominates
    bool physics_heavy = false;
    // two-level tree (chroma_geometry_create_instanced)
    bool instanced = false;
    InstView iv{};
    float4 *d_inst = nullptr;
    uint32_t *d_tri_base = nullptr, *d_tmpl_off_of_orig = nullptr;
    uint32_t *d_tmpl_colors = nullptr;
    // kept for the lazily assembled flat arrays of the GPUGeometry API (chroma_geometry_array)
    float *d_tmpl_vert = nullptr;
    uint32_t *d_tmpl_tri = nullptr, *d_tmpl_mcode = nullptr, *d_tmpl_nvert = nullptr, *d_vert_base = nullptr;
    void *d_tmpl_info = nullptr, *d_placements = nullptr;
    void *d_flat[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    uint64_t flat_nvertices = 0;
    uint32_t ntemplates = 0, inst_top_groups = 0, inst_top_leaves = 0;
    // 0: reference-format tree only (wide tree not built), 1: wide tree, 2: two-level tree
    int tree_kind = 0;
    // access-policy window last applied for this geometry (engine_wide.cu)
    int l2_window_mb = -1;
    cudaStream_t l2_window_stream = nullptr;
    // detector part
    bool has_detector = false;
    int32_t *d_solid_to_channel = nullptr;
    float *d_time_cdf_x = nullptr, *d_time_cdf_y = nullptr, *d_charge_cdf_x = nullptr,
          *d_charge_cdf_y = nullptr;
    uint64_t nsolids = 0;
    DetectorView det;
};

struct chroma_rng {
    chroma_ctx *ctx = nullptr;
    uint64_t seed = 0;
    uint32_t epoch = 0;  // bumped once per propagate / acquire / fill call
};

// per-call scratch of the wavefront propagate (engine_wavefront.cu), allocated on first use
struct WaveScratch {
    uint2 *aux = nullptr;     // per photon: x = leaf slot of last_hit_triangle, y = RNG words drawn
    int2 *hit = nullptr;      // per photon: x = hit leaf slot (-1 none, -2 NaN abort), y = distance bits
    uint32_t *queue[2] = {nullptr, nullptr};  // live photon ids, double buffered
    unsigned long long *ctr = nullptr;        // [0..3] queue lengths, [4..7] fetch heads (step mod 4)
    uint64_t capacity = 0;
};

// device copy of the host-side photon arrays (chroma.event.Photons layout)
struct StageView {
    const float *pos, *dir, *pol, *wl, *t, *weights;
    const int32_t *last_hit;
    const uint32_t *flags, *evidx;
    // value of a field whose array was not uploaded (pointer NULL): one vertex / start time for a whole
    // bomb, default weight 1, flags 0, last_hit -1, one event index per window
    float pos_v[3], t_v, weight_v;
    uint32_t flags_v, evidx_v;
    int32_t last_hit_v;
};
__host__ __device__ inline StageView stage_view_defaults() {
    StageView s{};
    s.weight_v = 1.0f;
    s.last_hit_v = -1;
    return s;
}

struct chroma_photons {
    chroma_ctx *ctx = nullptr;
    uint64_t capacity = 0, id_base = 0;
    PhotonView v{};
    bool has_channel = false;
    WaveScratch wave;
    // Staged upload (engine_photons.cu): the host arrays as they are (60 B/photon, field after
    // field) land in `stage` chunk by chunk on ctx->copy_stream; after each chunk the copy engine
    // writes the number of photons landed so far to *d_ready.  chroma_photons_propagate (wide
    // kernel) reads photons straight from the stage as they arrive; every other consumer calls
    // photons_join(), which waits for the transfer and packs the records.
    char *stage = nullptr;
    StageView sv{};
    uint64_t st_offset = 0, st_n = 0;
    unsigned long long *d_ready = nullptr;
    cudaEvent_t st_done = nullptr;
};

// launch parameters of the fused persistent propagate kernels (engine_wide.cu, engine_inst.cu)
struct WPropParams {
    uint64_t offset, n;
    uint64_t seed;
    uint32_t id_base, epoch;
    int max_steps, use_weights, scatter_first;
    int refill, iters, leaf_thresh;
    unsigned watchdog_spins;   // ~microseconds a warp waits for staged photons before it gives up (status bit 1)
    unsigned long long *head;
    uint32_t *nan_count, *status;
    // staged upload still in flight (engine_photons.cu): photons are read from the stage once
    // *ready (written by the copy engine) covers them; null = read the packed records
    StageView sv;
    uint64_t sv_offset;
    const unsigned long long *ready;
};

// wait for a staged upload, pack it into the records and release the stage (engine_photons.cu)
int photons_join(chroma_photons *p);
// release the stage of `p` on the context's stream without packing (the propagate kernel consumed it)
void photons_drop_stage(chroma_photons *p);
// host-side wait for the worker thread of a borrowed-source upload (engine_photons.cu); returns its status
int uploader_join(chroma_ctx *c);

// engine_core.cu: optics blob + tables of `d` into g->view.opt
int chroma_build_optics(chroma_ctx *c, chroma_geometry *g, const chroma_geometry_desc *d);

// engine_wide.cu
int chroma_wide_build(chroma_geometry *g, const uint32_t *d_reach, const uint32_t *d_leaf_flag,
                      const uint32_t *d_slot_of);
int chroma_wide_intersect(chroma_geometry *g, uint64_t n, const float *pos, const float *dir,
                          const int32_t *last_hit, int32_t *out_tri, float *out_dist,
                          uint32_t *out_counts);
int chroma_wide_propagate(chroma_photons *p, chroma_geometry *g, chroma_rng *rng, uint64_t offset,
                          uint64_t n, int max_steps, int use_weights, int scatter_first,
                          uint64_t *nan_aborted);

// engine_inst.cu: the same entry points on the two-level tree
int chroma_inst_intersect(chroma_geometry *g, uint64_t n, const float *pos, const float *dir,
                          const int32_t *last_hit, int32_t *out_tri, float *out_dist, uint32_t *out_counts);
int chroma_inst_prop_occupancy(chroma_ctx *c, bool ool, size_t *smem, int *occ);
void chroma_inst_prop_launch(chroma_geometry *g, chroma_photons *p, const WPropParams &prm, bool ool, bool staged,
                             unsigned grid, size_t smem);
// engine_wide_ool.cu / engine_inst_ool.cu: the same kernels built with the random draws out of line
int chroma_wide_ool_occupancy(chroma_ctx *c, size_t smem, int *occ);
void chroma_wide_ool_launch(chroma_geometry *g, chroma_photons *p, const WPropParams &prm, bool staged,
                            unsigned grid, size_t smem);
int chroma_inst_ool_occupancy(chroma_ctx *c, size_t smem, int *occ);
void chroma_inst_ool_launch(chroma_geometry *g, chroma_photons *p, const WPropParams &prm, bool staged, unsigned grid,
                            size_t smem);
int chroma_inst_flat_array(chroma_geometry *g, int which, void **dptr, uint64_t *count);

// 8-wide tree over items given by Morton codes + padded 16-bit boxes (engine_wide.cu); the caller
// owns out->groups and out->perm (cudaFree)
struct WideTreeOut {
    uint4 *groups = nullptr;
    uint32_t *perm = nullptr;   // sorted position -> input index
    uint32_t ngroups = 0, levels = 0, leaves = 0;
};
int wide_tree_from_boxes(chroma_ctx *c, uint64_t n, const uint64_t *d_codes, const uint32_t *d_boxes,
                         uint32_t leaf_max, WideTreeOut *out);

// engine_wavefront.cu
int chroma_wave_propagate(chroma_photons *p, chroma_geometry *g, chroma_rng *rng, uint64_t offset,
                          uint64_t n, int max_steps, int use_weights, int scatter_first,
                          uint64_t *nan_aborted);

struct chroma_daq {
    chroma_geometry *g = nullptr;
    chroma_ctx *ctx = nullptr;
    uint32_t ndaq = 1, nchannels = 0;
    uint32_t *d_time_int = nullptr, *d_q_int = nullptr, *d_hist = nullptr;
    float *d_time = nullptr, *d_q = nullptr;
};

struct chroma_pdf {
    chroma_ctx *ctx = nullptr;
    uint32_t nchannels = 0, tbins = 0, qbins = 0;
    float tmin = 0, tmax = 0, qmin = 0, qmax = 0;
    uint32_t *d_hitcount = nullptr, *d_pdf = nullptr;
};
