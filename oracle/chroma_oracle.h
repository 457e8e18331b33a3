/* chroma_oracle.h — CPU restatement of the reference's photon-propagation path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under chroma_b200/ may include, link or call this;
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm use it,
 * as the checker and the timed CPU baseline.
 *
 * Each function restates one reference routine (file:line cited in chroma_oracle.cpp) on the
 * reference's own data format: vertices float[nv][3], triangles uint32[nt][3],
 * material_codes uint32[nt], nodes uint32[nn][4] (packed uint4), flat optical tables.
 * The only substitution is the random generator: the reference draws from per-thread cuRAND
 * XORWOW states (device-only API); here, as in the engine, a Philox4x32-10 stream keyed by
 * (seed; photon id, call epoch, stream) is consumed in the reference's draw order.
 *
 * Parity pin: (1) the reference's own CUDA kernels run on the GPU (oracle/_ref/*.cubin,
 * tests/test_reference_gpu.py: bit-exact first hits); (2) the reference's own propagate.cu /
 * daq.cu / mesh.h compiled for the host and run here on the same Philox streams
 * (oracle/ref_cpu_wrapper.cpp -> oracle/_ref/libchroma_ref_cpu.so,
 * tests/test_oracle_vs_ref_cpu.py: identical first-hit triangles, photon histories and DAQ
 * output, distances to a few ulp); (3) node packing against the reference's test/test_bvh.py
 * vectors (tests/golden/).
 */
#ifndef CHROMA_ORACLE_H
#define CHROMA_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    const float *vertices; const uint32_t *triangles; const uint32_t *material_codes;
    const uint32_t *solid_id;                  /* may be NULL */
    uint64_t nvertices, ntriangles;
    const uint32_t *nodes; uint64_t nnodes;
    float world_origin[3]; float world_scale;
    uint32_t nwavelengths; float wavelength_start, wavelength_step;
    uint32_t ntimes; float time_start, time_step;
    uint32_t nmaterials; const float *material_props; const uint32_t *material_num_comp;
    uint32_t ncomp_total;
    const float *comp_reemission_prob, *comp_reemission_wvl_cdf, *comp_absorption_length,
                *comp_reemission_time_cdf;
    uint32_t nsurfaces; const uint8_t *surface_present; const float *surface_props;
    const uint32_t *surface_model, *surface_transmissive; const float *surface_thickness;
    const uint32_t *dichroic_nangles; const float *dichroic_angles, *dichroic_reflect,
                *dichroic_transmit;
    /* detector (may be all zero) */
    const int32_t *solid_id_to_channel_index; uint64_t nsolids; uint32_t nchannels;
    const float *time_cdf_x, *time_cdf_y; uint32_t time_cdf_len;
    const float *charge_cdf_x, *charge_cdf_y; uint32_t charge_cdf_len;
} oracle_geometry;

/* number of OpenMP threads the library will use (cores for cpu_baseline) */
int oracle_num_threads(void);
void oracle_set_num_threads(int n);

/* mesh.h:41-151: first hit per ray; out_counts[n][2] (may be NULL) = get_node calls and
 * triangle tests of mesh.h:65-66,75,84.  dir is normalised on entry (mesh.h:141).
 * fast_div != 0 reproduces --use_fast_math's reciprocal-multiply form of -o/d (not its
 * approximate reciprocal); 0 uses IEEE division. */
void oracle_intersect(const oracle_geometry *g, uint64_t n, const float *pos, const float *dir,
                      const int32_t *last_hit, int32_t *out_tri, float *out_dist,
                      uint32_t *out_counts);

/* photon bomb of chroma/benchmark.py:75-83 with chroma/sample.py:4-57 samplers on the Philox
 * stream (seed; id0+i, sequence, 3): dir/pol [n][3], wavelengths [n]; rot = 3x3 row major */
void oracle_generate_photons(uint64_t n, uint32_t id0, uint64_t seed, uint32_t sequence,
                             const float *rot, float cos_half, float wl_lo, float wl_hi,
                             float *dir, float *pol, float *wavelengths);

/* propagate.cu:217-319 + gpu/photon.py:193-259: run photons [0,n) for up to max_steps steps.
 * Arrays are event.Photons fields, updated in place.  id_base/epoch/seed key the RNG exactly
 * as the engine does.  out_steps (may be NULL): per-photon (steps, node visits, triangle
 * tests) accumulated over the call. */
void oracle_propagate(const oracle_geometry *g, uint64_t n, float *pos, float *dir, float *pol,
                      float *wavelengths, float *t, int32_t *last_hit, uint32_t *flags,
                      float *weights, uint64_t seed, uint32_t id_base, uint32_t epoch,
                      int max_steps, int use_weights, int scatter_first, uint64_t *out_steps);

/* daq.cu:35-86 run_daq for photons [0,n) + gpu/daq.py:54-99 begin/end: fills
 * t[nchannels], q[nchannels], hist[nchannels] */
void oracle_run_daq(const oracle_geometry *g, uint64_t n, const float *t, const uint32_t *flags,
                    const int32_t *last_hit, const float *weights, uint64_t seed,
                    uint32_t id_base, uint32_t epoch, float global_weight, float *out_t,
                    float *out_q, uint32_t *out_hist);
/* daq.cu:88-150 run_daq_many with ndaq replicas (out arrays hold ndaq * nchannels entries) */
void oracle_run_daq_many(const oracle_geometry *g, uint64_t n, const float *t, const uint32_t *flags,
                         const int32_t *last_hit, const float *weights, uint64_t seed, uint32_t id_base,
                         uint32_t epoch, float global_weight, int ndaq, int convert_all, float *out_t,
                         float *out_q, uint32_t *out_hist);

/* pdf.cu:9-32 bin_hits */
void oracle_bin_hits(int nchannels, const float *channel_q, const float *channel_time,
                     uint32_t *hitcount, int tbins, float tmin, float tmax, int qbins, float qmin,
                     float qmax, uint32_t *pdf);

/* bvh.cu:148-203 make_leaves (IEEE division) */
void oracle_make_leaves(const float *vertices, const uint32_t *triangles, uint64_t ntriangles,
                        const float world_origin[3], float world_scale, uint32_t *leaf_nodes,
                        uint64_t *morton_codes);

/* bvh/grid.py:11-95 make_recursive_grid_bvh with a STABLE sort of the Morton codes.
 * Returns the node count (or 0 on failure); nodes must hold 2*ntriangles+1024 entries,
 * layer_offsets 64; *nlayers receives the layer count. */
uint64_t oracle_build_bvh(const float *vertices, uint64_t nvertices, const uint32_t *triangles,
                          uint64_t ntriangles, int target_degree, uint32_t *nodes,
                          uint64_t *layer_offsets, uint32_t *nlayers, float world_origin[3],
                          float *world_scale);

/* the same from given leaf nodes [ntriangles][4] + 48-bit Morton codes (host part of grid.py) */
uint64_t oracle_build_bvh_from_leaves(const uint32_t *leaves, const uint64_t *codes,
                                      uint64_t ntriangles, int target_degree, uint32_t *nodes,
                                      uint64_t *layer_offsets, uint32_t *nlayers);

/* curand_uniform-style stream used by both sides: out[i] = i-th float of stream
 * (seed; id, epoch, stream) */
void oracle_uniform_stream(uint64_t seed, uint32_t id, uint32_t epoch, uint32_t stream, uint32_t n,
                           float *out);

#ifdef __cplusplus
}
#endif
#endif
