// engine_daq.cu — DAQ (channel time/charge building) and PDF binning.
//
// Replaces chroma/cuda/daq.cu + chroma/gpu/daq.py:37-99 (GPUDaq) and
// chroma/cuda/pdf.cu:9-32 + chroma/gpu/pdf.py:177-227 (GPUPDF.setup_pdf/add_hits_to_pdf).
#include "physics.cuh"

// daq.cu:5-20: times are ordered as raw uint32 bit patterns (valid for t >= 0 only — the
// reference's own quirk, kept: a negative sampled time never wins the atomicMin)
__device__ __forceinline__ unsigned float_to_sortable_int(float f) { return __float_as_uint(f); }

// daq.cu:35-86 run_daq
__global__ void run_daq_kernel(PhotonView v, uint64_t offset, uint64_t n, GeomView g,
                               DetectorView det, uint64_t nsolids, uint32_t detection_state,
                               uint64_t seed, uint32_t id_base, uint32_t epoch, float global_weight,
                               uint32_t *earliest_time_int, uint32_t *channel_q_int,
                               uint32_t *channel_histories) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        uint64_t pid = offset + i;
        uint4 st = v.st[pid];
        int triangle_id = (int)st.y;
        unsigned history = st.x;
        if (!(history & detection_state)) continue;
        uint32_t solid_id;
        if (!solid_of_triangle(g.tris, g.tri_slot, g.inst_tri_base, g.ninst, g.ntriangles, triangle_id, solid_id)) continue;
        if (solid_id >= nsolids) continue;
        int channel_index = det.solid_to_channel[solid_id];
        if (channel_index < 0) continue;

        const RngKey K = rng_key(seed, epoch, 1u);
        Rng rng;
        rng_init(rng, id_base + (uint32_t)pid);
        float weight = v.pw[pid].w * global_weight;
        if (rng_uniform(rng, K) < weight) {
            float time = v.pt[pid].w +
                         sample_cdf_tab(rng, K, det.time_cdf_len, det.time_cdf_x, det.time_cdf_y);
            unsigned time_int = float_to_sortable_int(time);
            float charge = sample_cdf_tab(rng, K, det.charge_cdf_len, det.charge_cdf_x, det.charge_cdf_y);
            unsigned charge_int = roundf(charge / det.charge_unit);
            atomicMin(earliest_time_int + channel_index, time_int);
            atomicAdd(channel_q_int + channel_index, charge_int);
            atomicOr(channel_histories + channel_index, history);
        }
    }
}

// daq.cu:88-150 run_daq_many: every detected photon is offered to `ndaq` DAQ replicas
__global__ void run_daq_many_kernel(PhotonView v, uint64_t offset, uint64_t n, GeomView g,
                                    DetectorView det, uint64_t nsolids, uint32_t detection_state,
                                    uint64_t seed, uint32_t id_base, uint32_t epoch,
                                    float global_weight, int ndaq, int channel_stride,
                                    uint32_t *earliest_time_int, uint32_t *channel_q_int,
                                    uint32_t *channel_histories) {
    // one warp per photon, lanes stride over the replicas
    const unsigned lane = threadIdx.x & 31;
    uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t i = warp; i < n; i += nwarps) {
        uint64_t pid = offset + i;
        uint4 st = v.st[pid];
        int triangle_id = (int)st.y;
        unsigned history = st.x;
        if (!(history & detection_state)) continue;
        uint32_t solid_id;
        if (!solid_of_triangle(g.tris, g.tri_slot, g.inst_tri_base, g.ninst, g.ntriangles, triangle_id, solid_id)) continue;
        if (solid_id >= nsolids) continue;
        int channel_index = det.solid_to_channel[solid_id];
        if (channel_index < 0) continue;
        float photon_time = v.pt[pid].w;
        float weight = v.pw[pid].w * global_weight;
        for (int k = lane; k < ndaq; k += 32) {
            const RngKey K = rng_key(seed, epoch, 0x10000u + (uint32_t)k);
            Rng rng;
            rng_init(rng, id_base + (uint32_t)pid);
            int channel_offset = channel_index + k * channel_stride;
            if (rng_uniform(rng, K) < weight) {
                // curand_normal: Box-Muller on two uniforms
                float u1 = rng_uniform(rng, K), u2 = rng_uniform(rng, K);
                float normal = sqrtf(-2.0f * logf(u1)) * sinf(2.0f * CHROMA_PI * u2);
                float time = photon_time + normal +
                             sample_cdf_tab(rng, K, det.time_cdf_len, det.time_cdf_x, det.time_cdf_y);
                unsigned time_int = float_to_sortable_int(time);
                float charge = sample_cdf_tab(rng, K, det.charge_cdf_len, det.charge_cdf_x, det.charge_cdf_y);
                unsigned charge_int = roundf(charge / det.charge_unit);
                atomicMin(earliest_time_int + channel_offset, time_int);
                atomicAdd(channel_q_int + channel_offset, charge_int);
                atomicOr(channel_histories + channel_offset, history);
            }
        }
    }
}

// gpu/daq.py:54-58: reset_earliest_time_int(1e9) + three fills, fused
__global__ void daq_begin_kernel(uint32_t n, uint32_t *time_int, uint32_t *q_int, float *q,
                                 uint32_t *hist) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        time_int[i] = __float_as_uint(1e9f);
        q_int[i] = 0;
        q[i] = 0.0f;
        hist[i] = 0;
    }
}

// daq.cu:152-172: convert_sortable_int_to_float + convert_charge_int_to_float, fused.
// The reference converts only the first nchannels charges (daq.cu:168); replicas beyond
// that keep q = 0 there.  All ndaq*nchannels entries are converted here.
__global__ void daq_end_kernel(uint32_t n, const uint32_t *time_int, const uint32_t *q_int,
                               float charge_unit, float *time, float *q) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        time[i] = __uint_as_float(time_int[i]);
        q[i] = q_int[i] * charge_unit;
    }
}

extern "C" int chroma_daq_create(chroma_geometry *g, uint32_t ndaq, chroma_daq **out) {
    ARG_CHECK(g && out && ndaq >= 1);
    if (!g->has_detector) {
        chroma_set_error("geometry has no detector attached (GPUDetector required)");
        return CHROMA_ERR_STATE;
    }
    chroma_ctx *c = g->ctx;
    CTX_BIND(c);
    chroma_daq *d = new chroma_daq();
    d->g = g;
    d->ctx = c;
    d->ndaq = ndaq;
    d->nchannels = (uint32_t)g->det.nchannels;
    size_t n = (size_t)d->nchannels * ndaq;
    size_t nb = (n ? n : 1) * 4;
    if (cudaMalloc(&d->d_time_int, nb) || cudaMalloc(&d->d_q_int, nb) || cudaMalloc(&d->d_hist, nb) ||
        cudaMalloc(&d->d_time, nb) || cudaMalloc(&d->d_q, nb)) {
        chroma_set_error("cudaMalloc daq: %s", cudaGetErrorString(cudaGetLastError()));
        chroma_daq_destroy(d);
        return CHROMA_ERR_CUDA;
    }
    cudaMemsetAsync(d->d_time_int, 0, nb, c->stream);
    cudaMemsetAsync(d->d_q_int, 0, nb, c->stream);
    cudaMemsetAsync(d->d_hist, 0, nb, c->stream);
    cudaMemsetAsync(d->d_time, 0, nb, c->stream);
    cudaMemsetAsync(d->d_q, 0, nb, c->stream);
    *out = d;
    return 0;
}

extern "C" int chroma_daq_destroy(chroma_daq *d) {
    if (!d) return 0;
    cudaSetDevice(d->ctx->device);
    cudaStreamSynchronize(d->ctx->stream);
    void *ptrs[] = {d->d_time_int, d->d_q_int, d->d_hist, d->d_time, d->d_q};
    for (void *p : ptrs)
        if (p) cudaFree(p);
    delete d;
    return 0;
}

extern "C" int chroma_daq_begin_acquire(chroma_daq *d) {
    ARG_CHECK(d);
    chroma_ctx *c = d->g->ctx;
    CTX_BIND(c);
    uint32_t n = d->nchannels * d->ndaq;
    if (n == 0) return 0;
    daq_begin_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(n, d->d_time_int, d->d_q_int, d->d_q, d->d_hist);
    KERNEL_CHECK(c);
    return 0;
}

extern "C" int chroma_daq_acquire(chroma_daq *d, chroma_photons *p, chroma_rng *rng,
                                  uint64_t offset, uint64_t n, float weight) {
    ARG_CHECK(d && p && rng);
    chroma_geometry *g = d->g;
    chroma_ctx *c = g->ctx;
    CTX_BIND(c);
    photons_join(p);
    ARG_CHECK(p->ctx == c && offset + n <= p->capacity);
    if (n == 0) return 0;
    uint32_t epoch = rng->epoch++;
    if (d->ndaq == 1) {
        run_daq_kernel<<<grid_for(n, 256, c, 8), 256, 0, c->stream>>>(
            p->v, offset, n, g->view, g->det, g->nsolids, (uint32_t)CHROMA_SURFACE_DETECT, rng->seed,
            (uint32_t)p->id_base, epoch, weight, d->d_time_int, d->d_q_int, d->d_hist);
    } else {
        run_daq_many_kernel<<<grid_for(n * 32, 256, c, 8), 256, 0, c->stream>>>(
            p->v, offset, n, g->view, g->det, g->nsolids, (uint32_t)CHROMA_SURFACE_DETECT, rng->seed,
            (uint32_t)p->id_base, epoch, weight, (int)d->ndaq, (int)d->nchannels, d->d_time_int,
            d->d_q_int, d->d_hist);
    }
    KERNEL_CHECK(c);
    return 0;
}

extern "C" int chroma_daq_end_acquire(chroma_daq *d) {
    ARG_CHECK(d);
    chroma_ctx *c = d->g->ctx;
    CTX_BIND(c);
    uint32_t n = d->nchannels * d->ndaq;
    if (n == 0) return 0;
    daq_end_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(n, d->d_time_int, d->d_q_int,
                                                           d->g->det.charge_unit, d->d_time, d->d_q);
    KERNEL_CHECK(c);
    return 0;
}

extern "C" int chroma_daq_get(chroma_daq *d, float *t, float *q, uint32_t *flags) {
    ARG_CHECK(d);
    chroma_ctx *c = d->g->ctx;
    CTX_BIND(c);
    size_t nb = (size_t)d->nchannels * d->ndaq * 4;
    if (t) CU_CHECK(cudaMemcpyAsync(t, d->d_time, nb, cudaMemcpyDeviceToHost, c->stream));
    if (q) CU_CHECK(cudaMemcpyAsync(q, d->d_q, nb, cudaMemcpyDeviceToHost, c->stream));
    if (flags) CU_CHECK(cudaMemcpyAsync(flags, d->d_hist, nb, cudaMemcpyDeviceToHost, c->stream));
    CU_CHECK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int chroma_daq_int_arrays(chroma_daq *d, void **time_int, void **q_int, uint64_t *count) {
    ARG_CHECK(d);
    if (time_int) *time_int = d->d_time_int;
    if (q_int) *q_int = d->d_q_int;
    if (count) *count = (uint64_t)d->nchannels * d->ndaq;
    return 0;
}

extern "C" int chroma_daq_arrays(chroma_daq *d, void **t, void **q, void **flags, uint64_t *count) {
    ARG_CHECK(d);
    if (t) *t = d->d_time;
    if (q) *q = d->d_q;
    if (flags) *flags = d->d_hist;
    if (count) *count = (uint64_t)d->nchannels * d->ndaq;
    return 0;
}

// ---------------------------------------------------------------------------------------
// One event over several GPUs: every rank runs run_daq on its share of the photons; between
// acquire and end_acquire the ranks' channel accumulators are merged.  The transport is the CALLER's
// collective (NCCL / MPI all-gather of the packed block; chroma_b200/dist.py uses torch.distributed on the
// context's stream); packing and the reduction are these two kernels.  The reference has no multi-GPU
// merge; the rule follows daq.cu:35-86: earliest time = min, charge = sum, history = or.
// ---------------------------------------------------------------------------------------
__global__ void daq_pack_kernel(const uint32_t *t, const uint32_t *q, const uint32_t *h, uint32_t n, uint32_t *out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = t[i]; out[n + i] = q[i]; out[2 * n + i] = h[i];
}

__global__ void daq_merge_kernel(uint32_t *t, uint32_t *q, uint32_t *h, uint32_t n, const uint32_t *gathered,
                                 uint32_t nranks) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // time bits of non-negative floats (and of the 1e9 "no hit" fill) order like the floats
    uint32_t tt = 0xFFFFFFFFu, qq = 0u, hh = 0u;
    for (uint32_t r = 0; r < nranks; r++) {
        const uint32_t *b = gathered + (size_t)r * 3 * n;
        tt = min(tt, b[i]);
        qq += b[n + i];
        hh |= b[2 * n + i];
    }
    t[i] = tt; q[i] = qq; h[i] = hh;
}

extern "C" int chroma_daq_pack(chroma_daq *d, uint32_t *out_device) {
    ARG_CHECK(d && out_device);
    chroma_ctx *c = d->g->ctx;
    CTX_BIND(c);
    const uint32_t n = d->nchannels * d->ndaq;
    if (n == 0) return 0;
    daq_pack_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(d->d_time_int, d->d_q_int, d->d_hist, n, out_device);
    KERNEL_CHECK(c);
    return 0;
}

extern "C" int chroma_daq_merge_gathered(chroma_daq *d, const uint32_t *gathered_device, uint32_t nranks) {
    ARG_CHECK(d && gathered_device && nranks >= 1);
    chroma_ctx *c = d->g->ctx;
    CTX_BIND(c);
    const uint32_t n = d->nchannels * d->ndaq;
    if (n == 0) return 0;
    daq_merge_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(d->d_time_int, d->d_q_int, d->d_hist, n, gathered_device, nranks);
    KERNEL_CHECK(c);
    return 0;
}

// ---------------------------------------------------------------------------------------
// PDF: pdf.cu:9-32 bin_hits
// ---------------------------------------------------------------------------------------
__global__ void bin_hits_kernel(int nchannels, const float *channel_q, const float *channel_time,
                                unsigned *hitcount, int tbins, float tmin, float tmax, int qbins,
                                float qmin, float qmax, unsigned *pdf) {
    int id = threadIdx.x + blockDim.x * blockIdx.x;
    if (id >= nchannels) return;
    unsigned int q = channel_q[id];  // float -> uint truncation, as the reference (pdf.cu:19)
    float t = channel_time[id];
    if (t < 1e8 && t >= tmin && t < tmax && q >= qmin && q < qmax) {
        hitcount[id] += 1;
        int tbin = (t - tmin) / (tmax - tmin) * tbins;
        int qbin = (q - qmin) / (qmax - qmin) * qbins;
        size_t bin = (size_t)id * ((size_t)tbins * qbins) + (size_t)tbin * qbins + qbin;
        pdf[bin] += 1;
    }
}

extern "C" int chroma_pdf_create(chroma_ctx *c, uint32_t nchannels, uint32_t tbins, float tmin,
                                 float tmax, uint32_t qbins, float qmin, float qmax, chroma_pdf **out) {
    CTX_BIND(c);
    ARG_CHECK(out && nchannels > 0 && tbins > 0 && qbins > 0);
    chroma_pdf *p = new chroma_pdf();
    p->ctx = c;
    p->nchannels = nchannels; p->tbins = tbins; p->qbins = qbins;
    p->tmin = tmin; p->tmax = tmax; p->qmin = qmin; p->qmax = qmax;
    size_t npdf = (size_t)nchannels * tbins * qbins;
    if (cudaMalloc(&p->d_hitcount, (size_t)nchannels * 4) || cudaMalloc(&p->d_pdf, npdf * 4)) {
        chroma_set_error("cudaMalloc pdf: %s", cudaGetErrorString(cudaGetLastError()));
        chroma_pdf_destroy(p);
        return CHROMA_ERR_CUDA;
    }
    *out = p;
    return chroma_pdf_clear(p);
}

extern "C" int chroma_pdf_destroy(chroma_pdf *p) {
    if (!p) return 0;
    cudaSetDevice(p->ctx->device);
    cudaStreamSynchronize(p->ctx->stream);
    if (p->d_hitcount) cudaFree(p->d_hitcount);
    if (p->d_pdf) cudaFree(p->d_pdf);
    delete p;
    return 0;
}

extern "C" int chroma_pdf_clear(chroma_pdf *p) {
    ARG_CHECK(p);
    chroma_ctx *c = p->ctx;
    CTX_BIND(c);
    CU_CHECK(cudaMemsetAsync(p->d_hitcount, 0, (size_t)p->nchannels * 4, c->stream));
    CU_CHECK(cudaMemsetAsync(p->d_pdf, 0, (size_t)p->nchannels * p->tbins * p->qbins * 4, c->stream));
    return 0;
}

extern "C" int chroma_pdf_add_device(chroma_pdf *p, const float *t_device, const float *q_device) {
    ARG_CHECK(p && t_device && q_device);
    chroma_ctx *c = p->ctx;
    CTX_BIND(c);
    bin_hits_kernel<<<(p->nchannels + 255) / 256, 256, 0, c->stream>>>(
        (int)p->nchannels, q_device, t_device, p->d_hitcount, (int)p->tbins, p->tmin, p->tmax,
        (int)p->qbins, p->qmin, p->qmax, p->d_pdf);
    KERNEL_CHECK(c);
    return 0;
}

extern "C" int chroma_pdf_add(chroma_pdf *p, chroma_daq *d) {
    ARG_CHECK(p && d);
    ARG_CHECK(d->nchannels == p->nchannels && d->g->ctx == p->ctx);
    return chroma_pdf_add_device(p, d->d_time, d->d_q);
}

extern "C" int chroma_pdf_get(chroma_pdf *p, uint32_t *hitcount, uint32_t *pdf) {
    ARG_CHECK(p);
    chroma_ctx *c = p->ctx;
    CTX_BIND(c);
    if (hitcount)
        CU_CHECK(cudaMemcpyAsync(hitcount, p->d_hitcount, (size_t)p->nchannels * 4, cudaMemcpyDeviceToHost, c->stream));
    if (pdf)
        CU_CHECK(cudaMemcpyAsync(pdf, p->d_pdf, (size_t)p->nchannels * p->tbins * p->qbins * 4, cudaMemcpyDeviceToHost, c->stream));
    CU_CHECK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int chroma_pdf_arrays(chroma_pdf *p, void **hitcount, void **pdf, uint64_t *pdf_count) {
    ARG_CHECK(p);
    if (hitcount) *hitcount = p->d_hitcount;
    if (pdf) *pdf = p->d_pdf;
    if (pdf_count) *pdf_count = (uint64_t)p->nchannels * p->tbins * p->qbins;
    return 0;
}
