/* chroma_b200.h — C-ABI of libchroma_b200.so (B200 / sm_100a photon-propagation engine).
 *
 * This is the drop-in boundary of the engine: it replaces the PyCUDA layer of the
 * reference (chroma/gpu/*.py launching chroma/cuda/*.cu).  Every entry point cites the
 * reference interface it stands in for (paths relative to the reference checkout).
 *
 * Conventions
 *   - plain C, opaque handles, caller-owned host buffers, library-owned device memory;
 *   - every function returns 0 on success, a negative chroma_status otherwise;
 *     chroma_last_error() returns a thread-local, human-readable message;
 *   - calls on one context are stream-ordered on the context's stream; functions that
 *     fill HOST buffers synchronise before returning, all others are asynchronous
 *     unless stated (mirrors pycuda's default-stream + Context.synchronize() usage,
 *     chroma/gpu/photon.py:112-114,256);
 *   - there is NO CPU fallback: without a CUDA device every compute call fails with
 *     CHROMA_ERR_CUDA.
 */
#ifndef CHROMA_B200_H
#define CHROMA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    CHROMA_OK = 0,
    CHROMA_ERR_CUDA = -1,      /* CUDA runtime error (text in chroma_last_error) */
    CHROMA_ERR_ARG = -2,       /* bad argument */
    CHROMA_ERR_STATE = -3,     /* call order / missing detector etc. */
    CHROMA_ERR_OVERFLOW = -4   /* traversal stack or node-buffer overflow */
} chroma_status;

/* photon history bits — chroma/cuda/photon.h:49-64, chroma/event.py:5-17 */
enum {
    CHROMA_NO_HIT = 0x1 << 0,
    CHROMA_BULK_ABSORB = 0x1 << 1,
    CHROMA_SURFACE_DETECT = 0x1 << 2,
    CHROMA_SURFACE_ABSORB = 0x1 << 3,
    CHROMA_RAYLEIGH_SCATTER = 0x1 << 4,
    CHROMA_REFLECT_DIFFUSE = 0x1 << 5,
    CHROMA_REFLECT_SPECULAR = 0x1 << 6,
    CHROMA_SURFACE_REEMIT = 0x1 << 7,
    CHROMA_SURFACE_TRANSMIT = 0x1 << 8,
    CHROMA_BULK_REEMIT = 0x1 << 9,
    CHROMA_CHERENKOV = 0x1 << 10,
    CHROMA_SCINTILLATION = 0x1 << 11,
    CHROMA_NAN_ABORT = (int)(0x1u << 31)
};

/* surface models — chroma/cuda/geometry_types.h:22 */
enum { CHROMA_SURFACE_DEFAULT = 0, CHROMA_SURFACE_COMPLEX = 1, CHROMA_SURFACE_WLS = 2,
       CHROMA_SURFACE_DICHROIC = 3 };

typedef struct chroma_ctx chroma_ctx;
typedef struct chroma_geometry chroma_geometry;
typedef struct chroma_rng chroma_rng;
typedef struct chroma_photons chroma_photons;
typedef struct chroma_daq chroma_daq;
typedef struct chroma_pdf chroma_pdf;
typedef struct chroma_bvh chroma_bvh;
typedef struct chroma_pdf_eval chroma_pdf_eval;
typedef struct chroma_kernel_pdf chroma_kernel_pdf;
typedef struct chroma_rays chroma_rays;

const char *chroma_last_error(void);
/* version string of the library, e.g. "chroma_b200 0.1 sm_100a" */
const char *chroma_version(void);

/* ---- context: chroma/gpu/tools.py:121-142 create_cuda_context ------------------------ */
int chroma_ctx_create(int device, chroma_ctx **out);
int chroma_ctx_destroy(chroma_ctx *ctx);
/* Tuning knobs (CHROMA_REFILL, CHROMA_WIDE_ITERS, CHROMA_WIDE_LEAF_THRESH, CHROMA_ISECT_REFILL,
 * CHROMA_ISECT_ITERS, CHROMA_ISECT_V, CHROMA_COUNT_WIDE, CHROMA_L2_WINDOW_MB, CHROMA_UPLOAD_DIRECT,
 * CHROMA_UPLOAD_CHUNKS, CHROMA_UPLOAD_THREADS, CHROMA_UPLOAD_WATCHDOG, CHROMA_RNG_OOL, CHROMA_PROP_V,
 * CHROMA_STRICT) are read from the environment once, in chroma_ctx_create; this re-reads them
 * (parameter sweeps).  Build-time settings of a geometry (CHROMA_WIDE_LEAF, CHROMA_WIDE_FAT,
 * CHROMA_INST_OPEN, CHROMA_OPTICS_SMEM_FLOATS) are read when the geometry is created. */
int chroma_ctx_reload_knobs(chroma_ctx *ctx);
int chroma_ctx_synchronize(chroma_ctx *ctx); /* Context.synchronize() */
/* the cudaStream_t every call on this context is ordered on: lets a caller put its own work (e.g. an
 * NCCL collective on the channel accumulators) in order with the library's without a host
 * synchronisation */
int chroma_ctx_stream(chroma_ctx *ctx, void **stream);
/* out[0]=free bytes, out[1]=total bytes, out[2]=SM count, out[3]=L2 bytes,
 * out[4]=max persisting L2 bytes, out[5]=compute capability major*10+minor */
int chroma_ctx_info(chroma_ctx *ctx, uint64_t out[6]);
/* number of engine kernels launched on this context since creation (bench.py gpu_launches) */
int chroma_ctx_launch_count(chroma_ctx *ctx, uint64_t *out);
/* CUDA-event timing on the context's stream: begin → work → end returns elapsed ms */
int chroma_ctx_timer_begin(chroma_ctx *ctx);
int chroma_ctx_timer_end(chroma_ctx *ctx, float *ms);
/* write `bytes` (> L2 size) of scratch to evict L2 between timed iterations */
int chroma_ctx_flush_l2(chroma_ctx *ctx);

/* ---- raw device buffers: stand-in for pycuda.gpuarray (to_gpu/get/fill) -------------- */
int chroma_malloc(chroma_ctx *ctx, size_t bytes, void **dptr);
int chroma_free(chroma_ctx *ctx, void *dptr);
int chroma_memcpy_h2d(chroma_ctx *ctx, void *dst, const void *src, size_t bytes);
int chroma_memcpy_d2h(chroma_ctx *ctx, void *dst, const void *src, size_t bytes); /* syncs */
int chroma_memcpy_d2d(chroma_ctx *ctx, void *dst, const void *src, size_t bytes);
int chroma_memset32(chroma_ctx *ctx, void *dst, uint32_t value, size_t count);
/* page-locked host memory for the end-to-end path */
int chroma_host_alloc(size_t bytes, void **hptr);
int chroma_host_free(void *hptr);

/* ---- geometry: chroma/gpu/geometry.py:13-259 GPUGeometry.__init__,
 *      structs chroma/cuda/geometry_types.h:4-82 ---------------------------------------- */
typedef struct {
    /* mesh (chroma/gpu/geometry.py:191-206) */
    const float *vertices;         /* [nvertices][3] */
    const uint32_t *triangles;     /* [ntriangles][3] */
    const uint32_t *material_codes;/* [ntriangles] inner<<24 | outer<<16 | surface<<8 */
    const uint32_t *colors;        /* [ntriangles] or NULL */
    const uint32_t *solid_id;      /* [ntriangles] or NULL */
    uint64_t nvertices, ntriangles;
    /* BVH in the reference's packed uint4 format (chroma/bvh/bvh.py:8-35); give either
     * host `nodes` or a device-resident `bvh` handle from chroma_bvh_build */
    const uint32_t *nodes;         /* [nnodes][4] or NULL */
    uint64_t nnodes;
    const chroma_bvh *bvh;         /* or NULL */
    float world_origin[3];
    float world_scale;
    /* uniform grids (chroma/gpu/geometry.py:15-30) */
    uint32_t nwavelengths; float wavelength_start, wavelength_step;
    uint32_t ntimes; float time_start, time_step;
    /* materials (chroma/gpu/geometry.py:47-100; Material struct geometry_types.h:4-20) */
    uint32_t nmaterials;
    const float *material_props;       /* [nmaterials][3][nwavelengths]: refractive_index,
                                          absorption_length, scattering_length */
    const uint32_t *material_num_comp; /* [nmaterials] */
    uint32_t ncomp_total;              /* sum(material_num_comp) */
    const float *comp_reemission_prob;     /* [ncomp_total][nwavelengths] */
    const float *comp_reemission_wvl_cdf;  /* [ncomp_total][nwavelengths] */
    const float *comp_absorption_length;   /* [ncomp_total][nwavelengths] */
    const float *comp_reemission_time_cdf; /* [ncomp_total][ntimes] */
    /* surfaces (chroma/gpu/geometry.py:108-186; Surface struct geometry_types.h:32-50) */
    uint32_t nsurfaces;
    const uint8_t *surface_present;    /* [nsurfaces] 0 = None entry */
    const float *surface_props;        /* [nsurfaces][8][nwavelengths]: detect, absorb, reemit,
                                          reflect_diffuse, reflect_specular, eta, k,
                                          reemission_cdf */
    const uint32_t *surface_model;     /* [nsurfaces] */
    const uint32_t *surface_transmissive; /* [nsurfaces] */
    const float *surface_thickness;    /* [nsurfaces] */
    /* dichroic (DichroicProps geometry_types.h:24-30) */
    const uint32_t *dichroic_nangles;  /* [nsurfaces] 0 = none; may be NULL */
    const float *dichroic_angles;      /* concatenated [sum nangles] */
    const float *dichroic_reflect;     /* concatenated [sum nangles][nwavelengths] */
    const float *dichroic_transmit;    /* concatenated [sum nangles][nwavelengths] */
} chroma_geometry_desc;

int chroma_geometry_create(chroma_ctx *ctx, const chroma_geometry_desc *desc,
                           chroma_geometry **out);
int chroma_geometry_destroy(chroma_geometry *g);
/* device-pointer getters for GPUGeometry's public attributes
 * (.nodes .vertices .triangles .material_codes .colors .solid_id_map):
 * which: 0 nodes(uint4) 1 vertices(float3) 2 triangles(uint3) 3 material_codes 4 colors
 *        5 solid_id_map; *count receives the element count */
int chroma_geometry_array(chroma_geometry *g, int which, void **dptr, uint64_t *count);
/* Frees the device copies of vertices, triangles, material codes and solid ids (5 GB at C3): the
 * engine searches its own triangle records and only hands these out as GPUGeometry attributes, which
 * a caller that still has the host mesh can upload itself when it wants them.  Colours stay (render
 * reads them).  chroma_geometry_array then fails for the freed arrays. */
int chroma_geometry_release_mesh(chroma_geometry *g);
/* out[0]=nnodes out[1]=ntriangles out[2]=nreachable_leaves out[3]=device bytes used
 * out[4]=optics floats staged in shared memory per CTA (0 = tables read from L1/L2)
 * out[5]=groups of the engine's wide search tree (0 = not built) out[6]=its depth out[7]=its leaf records */
int chroma_geometry_info(chroma_geometry *g, uint64_t out[8]);

/* detector: chroma/gpu/detector.py:14-39, struct chroma/cuda/detector.h:4-23 */
int chroma_detector_attach(chroma_geometry *g, const int32_t *solid_id_to_channel_index,
                           uint64_t nsolids, uint32_t nchannels,
                           const float *time_cdf_x, const float *time_cdf_y, uint32_t time_cdf_len,
                           const float *charge_cdf_x, const float *charge_cdf_y,
                           uint32_t charge_cdf_len);

/* ---- first-hit intersection: chroma/cuda/mesh.h:41-151 (intersect_mesh,
 *      distance_to_mesh); launch site chroma/benchmark.py:22-46 --------------------------
 * pos/dir: [n][3] float32 (dir is normalised on entry, mesh.h:141); last_hit: [n] or NULL;
 * out_triangle[n] (-1 = none), out_distance[n] (written for every ray; -1.0f when no hit).
 * `_device` takes device pointers (from chroma_malloc), `_host` takes host pointers and
 * performs the H2D/D2H copies itself (and synchronises). */
int chroma_intersect_device(chroma_geometry *g, uint64_t n, const float *pos, const float *dir,
                            const int32_t *last_hit, int32_t *out_triangle, float *out_distance);
int chroma_intersect_host(chroma_geometry *g, uint64_t n, const float *pos, const float *dir,
                          const int32_t *last_hit, int32_t *out_triangle, float *out_distance);
/* same traversal with the visit counters of chroma/cuda/mesh.h:65-66,75,84 switched on:
 * out_counts[n][2] = (get_node calls incl. root, triangle tests); device pointers */
int chroma_intersect_count_device(chroma_geometry *g, uint64_t n, const float *pos,
                                  const float *dir, const int32_t *last_hit,
                                  int32_t *out_triangle, float *out_distance,
                                  uint32_t *out_counts);
/* chroma_intersect_host for callers that may not allocate (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments and results as chroma_intersect_host (host
 * pointers, synchronises); the ray / result staging, the queue head and the status word live in one
 * file-scope device array of 256 + 36 B x CHROMA_STATIC_MAX_RAYS (9.0 MiB, held by every GPU that
 * loads the library until exit): the call performs no cudaMalloc / cudaFree / pool allocation.
 * n > CHROMA_STATIC_MAX_RAYS -> CHROMA_ERR_ARG; two-level geometries -> CHROMA_ERR_STATE.
 * All calls on a device share the array: calls that overlap on the GPU (two contexts or streams on
 * one device) race; calls on one context are stream-ordered and safe. */
#define CHROMA_STATIC_MAX_RAYS 262144u
int chroma_intersect_host_static(chroma_geometry *g, uint64_t n, const float *pos, const float *dir,
                                 const int32_t *last_hit, int32_t *out_triangle, float *out_distance);

/* ---- RNG: chroma/gpu/tools.py:75-84 get_rng_states / chroma/cuda/random.h:60-70 -------
 * The reference keeps `size` XORWOW states indexed by launch thread; the engine uses a
 * counter-based Philox4x32-10 keyed by (seed; photon id, call epoch), so `size` is
 * accepted and ignored. */
int chroma_rng_create(chroma_ctx *ctx, uint64_t size, uint64_t seed, chroma_rng **out);
int chroma_rng_destroy(chroma_rng *rng);
/* The call epoch is counter word 2 of every stream: each propagate / acquire / fill call uses the
 * current value and increments it.  Setting it keys the following calls by something the caller
 * chooses (e.g. 2 * global event number when events are dealt to several GPUs, so that the
 * result of an event does not depend on which rank ran it or on what that rank ran before). */
int chroma_rng_get_epoch(chroma_rng *rng, uint32_t *epoch);
int chroma_rng_set_epoch(chroma_rng *rng, uint32_t epoch);
/* fill out[n] (device pointer) with curand_uniform-style floats in (0,1]; test hook for
 * chroma/cuda/random.h:72-83 fill_uniform */
int chroma_rng_fill_uniform(chroma_rng *rng, uint64_t n, float *out_device);

/* ---- photons: chroma/gpu/photon.py:13-352 GPUPhotons ----------------------------------
 * Host layout is chroma.event.Photons: pos/dir/pol [n][3] float32, wavelengths/t/weights
 * float32[n], last_hit_triangles int32[n], flags/evidx uint32[n].  `id_base` offsets the
 * per-photon RNG key (global photon index of element 0) so results do not depend on how
 * photons are sharded over GPUs. */
int chroma_photons_create(chroma_ctx *ctx, uint64_t capacity, uint64_t id_base,
                          chroma_photons **out);
int chroma_photons_destroy(chroma_photons *p);
int chroma_photons_size(chroma_photons *p, uint64_t *capacity);
/* NULL last_hit → -1, NULL flags → 0, NULL weights → 1 (copy_triangles / copy_flags /
 * copy_weights = False, chroma/gpu/photon.py:36-45) */
int chroma_photons_upload(chroma_photons *p, uint64_t offset, uint64_t n, const float *pos,
                          const float *dir, const float *pol, const float *wavelengths,
                          const float *t, const int32_t *last_hit, const uint32_t *flags,
                          const float *weights, const uint32_t *evidx);
/* The same with fields given BY VALUE: where a pointer of `a` is NULL the value beside it stands for
 * the whole range (one vertex and start time for a photon bomb; weight 1, flags 0, last_hit -1 as
 * GPUPhotons(copy_weights/flags/triangles=False) does, chroma/gpu/photon.py:36-45; one event index per
 * window of a batch, chroma/sim.py:62-72) and nothing is uploaded for it.  dir, pol, wavelengths are
 * always arrays.  Pageable host arrays are copied through a page-locked ring owned by the context
 * (CHROMA_UPLOAD_THREADS host threads) and may be modified as soon as the call returns; page-locked
 * ones (chroma_host_alloc) are read by DMA until the next synchronising call. */
typedef struct chroma_photon_arrays {
    const float *pos, *dir, *pol, *wavelengths, *t, *weights;
    const int32_t *last_hit;
    const uint32_t *flags, *evidx;
    float pos_value[3], t_value, weight_value;
    uint32_t flags_value, evidx_value;
    int32_t last_hit_value;
    /* non-zero: the caller leaves PAGEABLE source arrays untouched until the next call that consumes
     * these photons has returned (chroma_photons_propagate, a download, chroma_ctx_synchronize ...); the
     * ring is then filled by a worker thread and this call returns at once, so a chroma_photons_propagate
     * issued next is already searching while the host is still copying */
    uint32_t borrow_sources;
} chroma_photon_arrays;
int chroma_photons_upload_arrays(chroma_photons *p, uint64_t offset, uint64_t n, const chroma_photon_arrays *a);
/* any output pointer may be NULL; synchronises */
int chroma_photons_download(chroma_photons *p, uint64_t offset, uint64_t n, float *pos,
                            float *dir, float *pol, float *wavelengths, float *t,
                            int32_t *last_hit, uint32_t *flags, float *weights,
                            uint32_t *evidx, int32_t *channel);
/* One field of photons [offset, offset+n) as a contiguous DEVICE array and back: what a caller that
 * hands `gpu_photons.pos` (a GPUArray in chroma/gpu/photon.py:29-46) to its own kernel binds to.
 * field: 0 pos, 1 dir, 2 pol (float[n][3]); 3 wavelengths, 4 t, 7 weights (float[n]);
 * 5 last_hit_triangles (int32[n]); 6 flags, 8 evidx, 9 channel (uint32[n]).  Stream-ordered. */
int chroma_photons_field_get(chroma_photons *p, uint64_t offset, uint64_t n, int field, void *out_device);
int chroma_photons_field_set(chroma_photons *p, uint64_t offset, uint64_t n, int field, const void *in_device);
/* device-to-device copy of a whole container (dst capacity >= src capacity) */
int chroma_photons_copy(chroma_photons *dst, chroma_photons *src);
/* photon_duplicate, chroma/cuda/propagate.cu:13-52: replicate [0,n) into copies 1..ncopies-1 */
int chroma_photons_duplicate(chroma_photons *p, uint64_t n, uint32_t ncopies);
/* ---- photon source on the device ----------------------------------
 * Stands in for the host-side construction of a photon bomb followed by its upload:
 * chroma/sample.py:4-30 uniform_sphere, chroma/sample.py:32-57 flashlight,
 * doc/source/example_detector.py:116-121 photon_bomb, chroma/benchmark.py:75-83 (the bomb of
 * the propagate benchmark) and chroma/gpu/photon.py:50-65 (the H2D copy it replaces).
 * Photon i of the call (global index g = id_base + offset + i) takes five uniforms in (0,1]
 * from the Philox stream (seed; g, sequence, stream 3), in this order:
 *   theta = 2*pi*u0;  u = cos_half_angle + u1*(1 - cos_half_angle);  c = sqrt(1 - u*u)
 *   dir   = rot * (c cos theta, c sin theta, u)          (cos_half_angle = -1: uniform_sphere;
 *                                                         cos(phi): flashlight(phi), rot = the
 *                                                         rotation of sample.py:38-45, row major)
 *   v     = uniform_sphere(u2, u3);  pol = normalize(cross(v, dir))   (benchmark.py:80)
 *   wavelength = wavelength_lo + u4*(wavelength_hi - wavelength_lo)   (benchmark.py:81)
 *   pos, t = t0, weight, evidx as given; flags = 0; last_hit_triangle = -1.
 * Results do not depend on the launch shape or on how [offset, offset+n) is split over calls
 * or GPUs.  Asynchronous on the context's stream. */
typedef struct chroma_source {
    float pos[3];
    float rot[9];
    float cos_half_angle;
    float wavelength_lo, wavelength_hi;
    float t0;
    float weight;
    uint32_t evidx;
    uint32_t sequence;
} chroma_source;
int chroma_photons_generate(chroma_photons *p, uint64_t offset, uint64_t n,
                            const chroma_source *src, uint64_t seed);
/* the step loop: chroma/gpu/photon.py:193-259 + chroma/cuda/propagate.cu:217-319.
 * Runs every photon in [offset, offset+n) for up to max_steps steps or until terminal;
 * may be called again to continue.  *nan_aborted (may be NULL) receives the number of
 * photons flagged NAN_ABORT in this call (synchronises when non-NULL). */
int chroma_photons_propagate(chroma_photons *p, chroma_geometry *g, chroma_rng *rng,
                             uint64_t offset, uint64_t n, int max_steps, int use_weights,
                             int scatter_first, uint64_t *nan_aborted);
/* count_photons/copy_photons chroma/cuda/propagate.cu:54-114 (GPUPhotons.select) */
int chroma_photons_select(chroma_photons *p, uint32_t target_flag, uint64_t offset, uint64_t n,
                          chroma_photons **out);
/* count_photon_hits/copy_photon_hits chroma/cuda/propagate.cu:147-214
 * (GPUPhotons.get_flat_hits): photons with flags&target_flag on a triangle whose solid maps
 * to a channel >= 0; the result carries a channel array (chroma_photons_download). */
int chroma_photons_flat_hits(chroma_photons *p, chroma_geometry *g, uint32_t target_flag,
                             uint64_t offset, uint64_t n, chroma_photons **out);
/* copy_photon_queue chroma/cuda/propagate.cu:116-144: gather photons by index */
int chroma_photons_gather(chroma_photons *p, const uint32_t *ids_host, uint64_t n,
                          chroma_photons **out);
/* ids (host, ascending) of photons in [offset,offset+n) that are not terminal — the
 * reference's output queue (chroma/cuda/propagate.cu:315-318); *count in/out capacity/result */
int chroma_photons_live_ids(chroma_photons *p, uint64_t offset, uint64_t n, uint32_t *ids_host,
                            uint64_t *count);

/* ---- DAQ: chroma/gpu/daq.py:37-99 GPUDaq, kernels chroma/cuda/daq.cu ------------------ */
int chroma_daq_create(chroma_geometry *detector, uint32_t ndaq, chroma_daq **out);
int chroma_daq_destroy(chroma_daq *d);
int chroma_daq_begin_acquire(chroma_daq *d); /* daq.cu:25-33 + fills, gpu/daq.py:54-58 */
/* run_daq (ndaq==1, daq.cu:35-86) or run_daq_many (daq.cu:88-150) */
int chroma_daq_acquire(chroma_daq *d, chroma_photons *p, chroma_rng *rng, uint64_t offset,
                       uint64_t n, float weight);
int chroma_daq_end_acquire(chroma_daq *d);   /* daq.cu:152-172 */
/* host copies of t (float), q (float), flags (uint32), each [nchannels*ndaq]; synchronises */
int chroma_daq_get(chroma_daq *d, float *t, float *q, uint32_t *flags);
/* device pointers of the integer accumulators (earliest_time_int, channel_q_int), for the
 * multi-GPU merge between acquire and end_acquire (min / sum over ranks) */
int chroma_daq_int_arrays(chroma_daq *d, void **time_int, void **q_int, uint64_t *count);
/* device pointers of the same arrays (GPUChannels.t/.q/.flags) */
int chroma_daq_arrays(chroma_daq *d, void **t, void **q, void **flags, uint64_t *count);
/* One event over several GPUs (the reference is single-GPU, the rule follows daq.cu:35-86):
 * between chroma_daq_acquire and chroma_daq_end_acquire every rank packs its channel accumulators
 * (uint32 [3][count]: earliest-time bits, charge ints, history words; count as chroma_daq_int_arrays reports
 * it), the CALLER moves the blocks with its own collective (an all-gather over NCCL or MPI into
 * [nranks][3][count]) and every rank reduces what it gathered: time = min, charge = sum, history = or.
 * Device pointers; both calls are ordered on the context's stream (chroma_ctx_stream), as the collective
 * must be.  chroma_b200/dist.py Communicator.reduce_channels is the Python caller. */
int chroma_daq_pack(chroma_daq *d, uint32_t *out_device);
int chroma_daq_merge_gathered(chroma_daq *d, const uint32_t *gathered_device, uint32_t nranks);

/* ---- PDF: chroma/gpu/pdf.py:177-227 GPUPDF, kernel chroma/cuda/pdf.cu:9-32 bin_hits ---- */
int chroma_pdf_create(chroma_ctx *ctx, uint32_t nchannels, uint32_t tbins, float tmin, float tmax,
                      uint32_t qbins, float qmin, float qmax, chroma_pdf **out);
int chroma_pdf_destroy(chroma_pdf *p);
int chroma_pdf_clear(chroma_pdf *p);
int chroma_pdf_add(chroma_pdf *p, chroma_daq *d); /* add_hits_to_pdf (DAQ replica 0) */
/* same with explicit device arrays t[nchannels], q[nchannels] (a GPUChannels slice) */
int chroma_pdf_add_device(chroma_pdf *p, const float *t_device, const float *q_device);
int chroma_pdf_get(chroma_pdf *p, uint32_t *hitcount, uint32_t *pdf); /* synchronises */
int chroma_pdf_arrays(chroma_pdf *p, void **hitcount, void **pdf, uint64_t *pdf_count);

/* ---- PDF evaluation at the observed hits: chroma/gpu/pdf.py:229-372 GPUPDF.setup_pdf_eval /
 *      clear_pdf_eval / accumulate_pdf_eval / get_pdf_eval, kernels chroma/cuda/pdf.cu:34-219
 *      (accumulate_bincount + accumulate_nearest_neighbor_block, fused into one launch).
 * event_hit / event_time: host arrays [nchannels] of the event of interest; accumulate() takes
 * GPUChannels.t of a GPUDaq with `ndaq` replicas (device, [ndaq][nchannels]).  get(): hitcount,
 * bincount uint32[nchannels]; nearest_mc float32[nhit][min_bin_content] (1e9 = empty), rows in
 * order of the hit channels; any output may be NULL. */
int chroma_pdf_eval_create(chroma_ctx *ctx, uint32_t nchannels, const uint32_t *event_hit,
                           const float *event_time, float min_twidth, float tmin, float tmax,
                           uint32_t min_bin_content, chroma_pdf_eval **out);
int chroma_pdf_eval_destroy(chroma_pdf_eval *e);
int chroma_pdf_eval_clear(chroma_pdf_eval *e);
int chroma_pdf_eval_accumulate(chroma_pdf_eval *e, const float *mc_time_device, uint32_t ndaq);
int chroma_pdf_eval_get(chroma_pdf_eval *e, uint32_t *hitcount, uint32_t *bincount,
                        float *nearest_mc, uint32_t *nhit); /* synchronises */

/* ---- kernel-estimated PDF: chroma/gpu/pdf.py:7-175 GPUKernelPDF, kernels
 *      chroma/cuda/pdf.cu:223-368 (accumulate_moments, accumulate_kernel_eval).  The bandwidth
 *      rule (compute_bandwidth, pdf.py:61-113) stays on the host, as in the reference. */
int chroma_kernel_pdf_create(chroma_ctx *ctx, uint32_t nchannels, float tmin, float tmax, float qmin,
                             float qmax, int time_only, chroma_kernel_pdf **out);
int chroma_kernel_pdf_destroy(chroma_kernel_pdf *k);
int chroma_kernel_pdf_clear_moments(chroma_kernel_pdf *k);
int chroma_kernel_pdf_accumulate_moments(chroma_kernel_pdf *k, const float *t_device,
                                         const float *q_device);
int chroma_kernel_pdf_get_moments(chroma_kernel_pdf *k, uint32_t *hitcount, float *tmom1, float *tmom2,
                                  float *qmom1, float *qmom2); /* synchronises */
int chroma_kernel_pdf_setup_kernel(chroma_kernel_pdf *k, const uint32_t *event_hit,
                                   const float *event_time, const float *event_charge,
                                   const float *inv_time_bandwidths,
                                   const float *inv_charge_bandwidths /* NULL = zeros */);
int chroma_kernel_pdf_clear_kernel(chroma_kernel_pdf *k);
int chroma_kernel_pdf_accumulate_kernel(chroma_kernel_pdf *k, const float *t_device,
                                        const float *q_device);
int chroma_kernel_pdf_get_kernel(chroma_kernel_pdf *k, uint32_t *hitcount, float *time_pdf_values,
                                 float *charge_pdf_values); /* synchronises */

/* ---- headless ray-cast render: chroma/gpu/render.py:8-73 GPURays, kernels
 *      chroma/cuda/render.cu:35-181 (render) and chroma/cuda/transform.cu:10-49.
 * pos / dir: host float32[n][3].  The per-ray hit lists (dx, color, dxlen: device arrays in the
 * reference's layout, [n][max_alpha_depth]) persist between renders for keep_last_render.
 * render(): pixels_device uint32[n] = a<<24 | r<<16 | g<<8 | b; alpha_depth <= max_alpha_depth. */
int chroma_rays_create(chroma_ctx *ctx, uint64_t n, const float *pos, const float *dir,
                       uint32_t max_alpha_depth, chroma_rays **out);
int chroma_rays_destroy(chroma_rays *r);
int chroma_rays_arrays(chroma_rays *r, void **pos, void **dir, void **dx, void **color, void **dxlen,
                       uint64_t *n);
int chroma_rays_translate(chroma_rays *r, const float v[3]);
int chroma_rays_rotate(chroma_rays *r, float phi, const float axis[3]);
int chroma_rays_rotate_around_point(chroma_rays *r, float phi, const float axis[3], const float point[3]);
int chroma_rays_render(chroma_rays *r, chroma_geometry *g, uint32_t alpha_depth, int keep_last_render,
                       uint32_t bg_color, uint32_t *pixels_device); /* synchronises */

/* ---- two-level ("instanced") geometry: SURVEY.md 8(f) rank 3 -------------------------------
 * Replaces the host flattening of every placed solid (chroma/geometry.py:360-367, placements from
 * chroma/detector.py:37-60 add_solid / add_pmt, 29,023 of them in chroma/demo/__init__.py:32-60)
 * and the upload of the flattened mesh + BVH (chroma/gpu/geometry.py:191-255): every distinct solid
 * ("template") is handed over once in its own frame, every placement as (template, rotation,
 * displacement); the device makes the boxes, one 8-wide tree per template and a top tree over the
 * placements.  Triangle ids and solid ids are those of the flattened mesh (solid = placement in
 * add order, triangles of solid i follow those of solid i-1).  Statistical tier: the triangle test
 * runs in the template frame, so distances differ from the flat mesh in the last bits. */
typedef struct chroma_instances_desc {
    uint32_t ntemplates;
    const float *vertices;             /* [vertex_offset[ntemplates]][3], each template in its own frame */
    const uint32_t *triangles;         /* [triangle_offset[ntemplates]][3], vertex indices inside the template */
    const uint32_t *material_codes;    /* [triangle_offset[ntemplates]] inner<<24 | outer<<16 | surface<<8 */
    const uint32_t *colors;            /* same length, or NULL */
    const uint64_t *vertex_offset;     /* [ntemplates + 1] */
    const uint64_t *triangle_offset;   /* [ntemplates + 1] */
    uint64_t ninstances;               /* placements in add order (= solid ids) */
    const uint32_t *instance_template; /* [ninstances] */
    const float *rotations;            /* [ninstances][9] row major: world = R v + d (np.inner(v, R) + d) */
    const float *displacements;        /* [ninstances][3] */
} chroma_instances_desc;
/* `optics`: only the wavelength / time grids and the material / surface tables are read */
int chroma_geometry_create_instanced(chroma_ctx *ctx, const chroma_geometry_desc *optics,
                                     const chroma_instances_desc *inst, chroma_geometry **out);
/* which search structure is live: out[0] = 0 reference-format tree only (the wide tree could not be
 * built: half the rate), 1 wide tree (default), 2 two-level tree; out[1] = placements, out[2] =
 * templates, out[3] = triangle slots held on the device, out[4] = groups of the top tree,
 * out[5] = all groups, out[6] = bytes of the search structure (groups + triangle and placement
 * records), out[7] = leaves of the top tree (placements, big ones split into patches).  CHROMA_STRICT=1 makes chroma_geometry_create fail instead of falling back
 * to the reference-format tree. */
int chroma_geometry_tree(chroma_geometry *g, uint64_t out[8]);

/* ---- BVH construction on device: chroma/bvh/grid.py:11-95 make_recursive_grid_bvh,
 *      kernels chroma/cuda/bvh.cu:148-203,269-308,364-384,530-543 ----------------------- */
int chroma_bvh_build(chroma_ctx *ctx, const float *vertices, uint64_t nvertices,
                     const uint32_t *triangles, uint64_t ntriangles, int target_degree,
                     chroma_bvh **out);
/* the host part of grid.py:24-95 (sort, grouping, parents, concatenation, chain collapse) on
 * ready-made leaves uint32[n][4] + Morton codes, as chroma/bvh/grid.py receives them from
 * create_leaf_nodes (chroma/gpu/bvh.py:17-78); ties between equal codes keep input order */
int chroma_bvh_build_from_leaves(chroma_ctx *ctx, const uint32_t *leaves, const uint64_t *codes,
                                 uint64_t nleaves, const float world_origin[3], float world_scale,
                                 int target_degree, chroma_bvh **out);
int chroma_bvh_destroy(chroma_bvh *b);
/* out[0]=nnodes out[1]=nlayers */
int chroma_bvh_info(chroma_bvh *b, uint64_t out[2], float world_origin[3], float *world_scale);
/* nodes[nnodes][4] uint32 (packed uint4), layer_offsets[nlayers] */
int chroma_bvh_get(chroma_bvh *b, uint32_t *nodes, uint64_t *layer_offsets);
/* building blocks kept for chroma/gpu/bvh.py API parity (host buffers in and out) */
int chroma_bvh_make_leaves(chroma_ctx *ctx, const float *vertices, uint64_t nvertices,
                           const uint32_t *triangles, uint64_t ntriangles,
                           const float world_origin[3], float world_scale,
                           uint32_t *leaf_nodes, uint64_t *morton_codes);
int chroma_bvh_merge_nodes_detailed(chroma_ctx *ctx, const uint32_t *nodes, uint64_t nnodes,
                                    const int32_t *first_child, const int32_t *nchild,
                                    uint64_t nparent, uint32_t *parent_nodes);
int chroma_bvh_collapse_chains(chroma_ctx *ctx, uint32_t *nodes, uint64_t nnodes,
                               const uint64_t *layer_bounds, uint64_t nbounds);

#ifdef __cplusplus
}
#endif
#endif /* CHROMA_B200_H */
