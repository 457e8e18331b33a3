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
#include "shim.h"

// ---- the reference's own sources ---------------------------------------------------------
#include "propagate.cu"   // chroma/cuda/propagate.cu
// run_daq_many (daq.cu:88-150) shares per-photon values between the threads of a block through
// __shared__ variables set by thread 0: the "threads" of one block are run one after the other by
// the same host thread below, so a thread-local static keeps what thread 0 wrote.
#undef __shared__
#define __shared__ static thread_local
#include "daq.cu"         // chroma/cuda/daq.cu
#undef __shared__
#define __shared__
// -------------------------------------------------------------------------------------------

#include "ref_cpu/geometry.h"


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


/* chroma/cuda/daq.cu:88-150 run_daq_many with the launch shape of chroma/gpu/daq.py:78-89 (one block
 * per photon); blockDim.x = ndaq here so that thread i serves replica i alone and its stream can be
 * keyed like the engine's (seed; photon id, epoch, 0x10000 + i).  convert_all = 0 keeps the
 * reference's end_acquire (daq.cu:161-172 converts the first nchannels charges only). */
extern "C" void ref_cpu_run_daq_many(const oracle_geometry *o, uint64_t n, float *t, uint32_t *flags,
                                     int32_t *last_hit, float *weights, uint64_t seed, uint32_t id_base,
                                     uint32_t epoch, float global_weight, int ndaq, int convert_all,
                                     float *out_t, float *out_q, uint32_t *out_hist) {
    RefGeometry R(o);
    const int nch = R.det.nchannels;
    const int tot = nch * ndaq;
    std::vector<unsigned int> time_int(tot), q_int(tot, 0u);
    blockDim.x = 1; threadIdx.x = 0;
    for (int i = 0; i < tot; i++) { blockIdx.x = i; reset_earliest_time_int(1e9f, tot, time_int.data()); }
    for (int i = 0; i < tot; i++) { out_hist[i] = 0u; out_q[i] = 0.0f; }
    std::vector<curandState> states((size_t)ndaq);
    for (uint64_t ph = 0; ph < n; ph++) {
        for (int k = 0; k < ndaq; k++) refcpu_rng_seed(&states[k], seed, id_base + (uint32_t)ph, epoch, 0x10000u + (uint32_t)k);
        // s[id] with id = threadIdx.x + blockDim.x * blockIdx.x: hand the kernel a base pointer such
        // that this block's threads land on states[0..ndaq)
        curandState *base = states.data() - (size_t)ndaq * ph;
        for (int k = 0; k < ndaq; k++) {
            blockIdx.x = (unsigned)ph; blockDim.x = (unsigned)ndaq; threadIdx.x = (unsigned)k;
            run_daq_many(base, 0x1 << 2, 0, (int)n, t, flags, last_hit, weights, (int *)o->solid_id, &R.det,
                         time_int.data(), q_int.data(), out_hist, ndaq, nch, global_weight);
        }
    }
    blockDim.x = 1; threadIdx.x = 0;
    for (int i = 0; i < tot; i++) {
        blockIdx.x = i;
        convert_sortable_int_to_float(tot, time_int.data(), out_t);
        convert_charge_int_to_float(&R.det, q_int.data(), out_q);
    }
    if (convert_all)
        for (int i = nch; i < tot; i++) out_q[i] = q_int[i] * R.det.charge_unit;
}
