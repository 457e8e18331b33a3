// engine_pdfeval.cu — likelihood PDF evaluation at the observed hit times.
//
// Replaces chroma/cuda/pdf.cu:34-368 (accumulate_bincount, accumulate_nearest_neighbor[_block],
// accumulate_moments, accumulate_kernel_eval) and their hosts chroma/gpu/pdf.py:7-175 (GPUKernelPDF)
// and :229-372 (GPUPDF.setup_pdf_eval / accumulate_pdf_eval / get_pdf_eval).
//
// The reference runs the variable-bin evaluation as two launches with a host synchronise in
// between: accumulate_bincount fills per-hit work queues in global memory, then one block per hit
// channel merges its queue into the sorted nearest-neighbour list with an insertion sort done by
// thread 0.  Here one launch does both: a thread owns a channel, walks the DAQ replicas in order
// (the running bin count decides which distances are kept, pdf.cu:83-87, so the order matters) and
// inserts each kept distance straight into its sorted list.  The list a channel ends with is the
// k smallest of (valid old entries + kept distances), exactly what piksrt + truncate produce.
#include "common.cuh"

struct chroma_pdf_eval {
    chroma_ctx *ctx = nullptr;
    uint32_t nchannels = 0, nhit = 0, min_bin_content = 0;
    float min_twidth = 0, tmin = 0, tmax = 0;
    uint32_t *d_event_hit = nullptr, *d_hit_offset = nullptr, *d_hitcount = nullptr, *d_bincount = nullptr;
    float *d_event_time = nullptr, *d_nearest = nullptr;
};

struct chroma_kernel_pdf {
    chroma_ctx *ctx = nullptr;
    uint32_t nchannels = 0;
    int time_only = 1;
    float tmin = 0, tmax = 0, qmin = 0, qmax = 0;
    uint32_t *d_hitcount = nullptr, *d_event_hit = nullptr;
    float *d_tmom1 = nullptr, *d_tmom2 = nullptr, *d_qmom1 = nullptr, *d_qmom2 = nullptr;
    float *d_event_time = nullptr, *d_event_charge = nullptr, *d_inv_tbw = nullptr, *d_inv_qbw = nullptr;
    float *d_time_pdf = nullptr, *d_charge_pdf = nullptr;
    bool kernel_ready = false;
};

// pdf.cu:34-149 in one pass.  mc_time is [ndaq][nchannels] (GPUChannels.t of a GPUDaq with ndaq replicas)
__global__ void pdf_eval_accumulate_kernel(int nchannels, int ndaq, const unsigned *event_hit,
                                           const float *event_time, const float *mc_time,
                                           unsigned *hitcount, unsigned *bincount, float min_twidth,
                                           float tmin, float tmax, int min_bin_content,
                                           const unsigned *hit_offset, float *nearest_mc) {
    int ch = threadIdx.x + blockDim.x * blockIdx.x;
    if (ch >= nchannels) return;
    // the reference keeps both counters in float while it loops (pdf.cu:52-53)
    float ch_hitcount = hitcount[ch];
    float ch_bincount = bincount[ch];
    const float ev_time = event_time[ch];
    const int ev_hit = event_hit[ch];
    float *list = nearest_mc + (size_t)min_bin_content * hit_offset[ch];
    int len = -1;  // valid prefix of the list, counted lazily (pdf.cu:121-129: up to the first entry > 1e8)
    for (int i = 0; i < ndaq; i++) {
        float t = mc_time[(size_t)nchannels * i + ch];
        if (t >= 1e8f) continue;
        if (t < tmin || t > tmax) continue;
        ch_hitcount += 1;
        if (!ev_hit) continue;
        float distance = fabsf(t - ev_time);
        if (distance < min_twidth * 0.5f) ch_bincount += 1;
        if (ch_bincount < min_bin_content) {
            if (len < 0) {
                len = 0;
                while (len < min_bin_content && !(list[len] > 1e8f)) len++;
            }
            // insert into the sorted prefix; when the list is full the largest entry falls off
            int pos = len < min_bin_content ? len : min_bin_content - 1;
            if (len == min_bin_content && !(distance < list[pos])) continue;
            while (pos > 0 && list[pos - 1] > distance) {
                list[pos] = list[pos - 1];
                pos--;
            }
            list[pos] = distance;
            if (len < min_bin_content) len++;
        }
    }
    hitcount[ch] = ch_hitcount;
    bincount[ch] = ch_bincount;
}

// pdf.cu:223-266
__global__ void kpdf_moments_kernel(int time_only, int nchannels, const float *mc_time,
                                    const float *mc_charge, float tmin, float tmax, float qmin,
                                    float qmax, unsigned *mom0, float *t_mom1, float *t_mom2,
                                    float *q_mom1, float *q_mom2) {
    int id = threadIdx.x + blockDim.x * blockIdx.x;
    if (id >= nchannels) return;
    float t = mc_time[id];
    if (time_only) {
        if (t < tmin || t > tmax) return;
        mom0[id] += 1;
        t_mom1[id] += t;
        t_mom2[id] += t * t;
    } else {
        float q = mc_charge[id];
        if (t < tmin || t > tmax || q < qmin || q > qmax) return;
        mom0[id] += 1;
        t_mom1[id] += t;
        t_mom2[id] += t * t;
        q_mom1[id] += q;
        q_mom2[id] += q * q;
    }
}

// pdf.cu:268-368
__global__ void kpdf_eval_kernel(int time_only, int nchannels, const unsigned *event_hit,
                                 const float *event_time, const float *event_charge,
                                 const float *mc_time, const float *mc_charge, float tmin, float tmax,
                                 float qmin, float qmax, const float *inv_time_bandwidths,
                                 const float *inv_charge_bandwidths, unsigned *hitcount,
                                 float *time_pdf_values, float *charge_pdf_values) {
    const float invroot2 = 0.70710678118654746f;
    const float rootPiBy2 = 1.2533141373155001f;
    int id = threadIdx.x + blockDim.x * blockIdx.x;
    if (id >= nchannels) return;
    float t = mc_time[id];
    if (time_only) {
        if (t < tmin || t > tmax) return;
        hitcount[id] += 1;
        if (!event_hit[id]) return;
        float inv_bandwidth = inv_time_bandwidths[id];
        float arg = (t - event_time[id]) * inv_bandwidth;
        float term = expf(-0.5f * arg * arg) * inv_bandwidth;
        float norm = tmax - tmin;
        if (inv_bandwidth > 0.0f) {
            float loarg = (tmin - t) * inv_bandwidth * invroot2;
            float hiarg = (tmax - t) * inv_bandwidth * invroot2;
            norm = (erff(hiarg) - erff(loarg)) * rootPiBy2;
        }
        time_pdf_values[id] += term / norm;
    } else {
        float q = mc_charge[id];
        if (t < tmin || t > tmax || q < qmin || q > qmax) return;
        hitcount[id] += 1;
        if (!event_hit[id]) return;
        float inv_bandwidth = inv_time_bandwidths[id];
        float arg = (t - event_time[id]) * inv_bandwidth;
        float norm = tmax - tmin;
        if (inv_bandwidth > 0.0f) {
            float loarg = (tmin - t) * inv_bandwidth * invroot2;
            float hiarg = (tmax - t) * inv_bandwidth * invroot2;
            norm = (erff(hiarg) - erff(loarg)) * rootPiBy2;
        }
        float term = expf(-0.5f * arg * arg);
        time_pdf_values[id] += term / norm;
        inv_bandwidth = inv_charge_bandwidths[id];
        arg = (q - event_charge[id]) * inv_bandwidth;
        norm = qmax - qmin;
        if (inv_bandwidth > 0.0f) {
            float loarg = (qmin - q) * inv_bandwidth * invroot2;
            float hiarg = (qmax - q) * inv_bandwidth * invroot2;
            norm = (erff(hiarg) - erff(loarg)) * rootPiBy2;
        }
        term = expf(-0.5f * arg * arg);
        charge_pdf_values[id] += term / norm;
    }
}

__global__ void fill_f32_kernel(float *dst, float value, size_t n) {
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = value;
}

template <typename T>
static int dev_alloc(T **p, size_t count) {
    CU_CHECK(cudaMalloc((void **)p, (count ? count : 1) * sizeof(T)));
    return 0;
}
template <typename T>
static int dev_upload(chroma_ctx *c, T **p, const T *src, size_t count) {
    int rc = dev_alloc(p, count);
    if (rc) return rc;
    if (count) CU_CHECK(cudaMemcpyAsync(*p, src, count * sizeof(T), cudaMemcpyHostToDevice, c->stream));
    return 0;
}
#define TRYE(x)                                                                            \
    do {                                                                                   \
        int _r = (x);                                                                      \
        if (_r) { cleanup(); return _r; }                                                  \
    } while (0)

// ---- GPUPDF.setup_pdf_eval / clear_pdf_eval / accumulate_pdf_eval / get_pdf_eval -------------
extern "C" int chroma_pdf_eval_destroy(chroma_pdf_eval *e) {
    if (!e) return 0;
    cudaSetDevice(e->ctx->device);
    cudaStreamSynchronize(e->ctx->stream);
    void *ptrs[] = {e->d_event_hit, e->d_hit_offset, e->d_hitcount, e->d_bincount, e->d_event_time, e->d_nearest};
    for (void *p : ptrs) if (p) cudaFree(p);
    delete e;
    return 0;
}

extern "C" int chroma_pdf_eval_clear(chroma_pdf_eval *e) {
    ARG_CHECK(e);
    chroma_ctx *c = e->ctx;
    CTX_BIND(c);
    CU_CHECK(cudaMemsetAsync(e->d_hitcount, 0, (size_t)e->nchannels * 4, c->stream));
    CU_CHECK(cudaMemsetAsync(e->d_bincount, 0, (size_t)e->nchannels * 4, c->stream));
    size_t n = (size_t)e->nhit * e->min_bin_content;
    if (n) {
        fill_f32_kernel<<<grid_for(n, 256, c, 4), 256, 0, c->stream>>>(e->d_nearest, 1e9f, n);
        KERNEL_CHECK(c);
    }
    return 0;
}

extern "C" int chroma_pdf_eval_create(chroma_ctx *c, uint32_t nchannels, const uint32_t *event_hit,
                                      const float *event_time, float min_twidth, float tmin, float tmax,
                                      uint32_t min_bin_content, chroma_pdf_eval **out) {
    CTX_BIND(c);
    ARG_CHECK(out && nchannels > 0 && event_hit && event_time && min_bin_content > 0);
    chroma_pdf_eval *e = new chroma_pdf_eval();
    auto cleanup = [&]() { chroma_pdf_eval_destroy(e); };
    e->ctx = c;
    e->nchannels = nchannels;
    e->min_bin_content = min_bin_content;
    e->min_twidth = min_twidth; e->tmin = tmin; e->tmax = tmax;
    // chroma/gpu/pdf.py:266-270: hit offset = max(0, cumsum(hit) - 1)
    std::vector<uint32_t> hit01(nchannels), off(nchannels);
    uint32_t run = 0;
    for (uint32_t i = 0; i < nchannels; i++) {
        hit01[i] = event_hit[i] ? 1u : 0u;
        run += hit01[i];
        off[i] = run ? run - 1 : 0;
    }
    e->nhit = run;
    TRYE(dev_upload(c, &e->d_event_hit, hit01.data(), nchannels));
    TRYE(dev_upload(c, &e->d_hit_offset, off.data(), nchannels));
    TRYE(dev_upload(c, &e->d_event_time, event_time, nchannels));
    TRYE(dev_alloc(&e->d_hitcount, nchannels));
    TRYE(dev_alloc(&e->d_bincount, nchannels));
    TRYE(dev_alloc(&e->d_nearest, (size_t)(e->nhit ? e->nhit : 1) * min_bin_content));
    TRYE(chroma_pdf_eval_clear(e));
    if (cudaStreamSynchronize(c->stream) != cudaSuccess) {   // host vectors go out of scope
        chroma_set_error("pdf_eval setup: %s", cudaGetErrorString(cudaGetLastError()));
        cleanup();
        return CHROMA_ERR_CUDA;
    }
    *out = e;
    return 0;
}

extern "C" int chroma_pdf_eval_accumulate(chroma_pdf_eval *e, const float *mc_time_device, uint32_t ndaq) {
    ARG_CHECK(e && mc_time_device && ndaq > 0);
    chroma_ctx *c = e->ctx;
    CTX_BIND(c);
    pdf_eval_accumulate_kernel<<<(e->nchannels + 127) / 128, 128, 0, c->stream>>>(
        (int)e->nchannels, (int)ndaq, e->d_event_hit, e->d_event_time, mc_time_device, e->d_hitcount,
        e->d_bincount, e->min_twidth, e->tmin, e->tmax, (int)e->min_bin_content, e->d_hit_offset, e->d_nearest);
    KERNEL_CHECK(c);
    return 0;
}

extern "C" int chroma_pdf_eval_get(chroma_pdf_eval *e, uint32_t *hitcount, uint32_t *bincount,
                                   float *nearest_mc, uint32_t *nhit) {
    ARG_CHECK(e);
    chroma_ctx *c = e->ctx;
    CTX_BIND(c);
    if (hitcount) CU_CHECK(cudaMemcpyAsync(hitcount, e->d_hitcount, (size_t)e->nchannels * 4, cudaMemcpyDeviceToHost, c->stream));
    if (bincount) CU_CHECK(cudaMemcpyAsync(bincount, e->d_bincount, (size_t)e->nchannels * 4, cudaMemcpyDeviceToHost, c->stream));
    if (nearest_mc && e->nhit)
        CU_CHECK(cudaMemcpyAsync(nearest_mc, e->d_nearest, (size_t)e->nhit * e->min_bin_content * 4, cudaMemcpyDeviceToHost, c->stream));
    CU_CHECK(cudaStreamSynchronize(c->stream));
    if (nhit) *nhit = e->nhit;
    return 0;
}

// ---- GPUKernelPDF ------------------------------------------------------------------------------
extern "C" int chroma_kernel_pdf_destroy(chroma_kernel_pdf *k) {
    if (!k) return 0;
    cudaSetDevice(k->ctx->device);
    cudaStreamSynchronize(k->ctx->stream);
    void *ptrs[] = {k->d_hitcount, k->d_event_hit, k->d_tmom1, k->d_tmom2, k->d_qmom1, k->d_qmom2,
                    k->d_event_time, k->d_event_charge, k->d_inv_tbw, k->d_inv_qbw, k->d_time_pdf, k->d_charge_pdf};
    for (void *p : ptrs) if (p) cudaFree(p);
    delete k;
    return 0;
}

extern "C" int chroma_kernel_pdf_clear_moments(chroma_kernel_pdf *k) {
    ARG_CHECK(k);
    chroma_ctx *c = k->ctx;
    CTX_BIND(c);
    size_t nb = (size_t)k->nchannels * 4;
    CU_CHECK(cudaMemsetAsync(k->d_hitcount, 0, nb, c->stream));
    CU_CHECK(cudaMemsetAsync(k->d_tmom1, 0, nb, c->stream));
    CU_CHECK(cudaMemsetAsync(k->d_tmom2, 0, nb, c->stream));
    CU_CHECK(cudaMemsetAsync(k->d_qmom1, 0, nb, c->stream));
    CU_CHECK(cudaMemsetAsync(k->d_qmom2, 0, nb, c->stream));
    return 0;
}

extern "C" int chroma_kernel_pdf_create(chroma_ctx *c, uint32_t nchannels, float tmin, float tmax,
                                        float qmin, float qmax, int time_only, chroma_kernel_pdf **out) {
    CTX_BIND(c);
    ARG_CHECK(out && nchannels > 0);
    chroma_kernel_pdf *k = new chroma_kernel_pdf();
    auto cleanup = [&]() { chroma_kernel_pdf_destroy(k); };
    k->ctx = c;
    k->nchannels = nchannels;
    k->tmin = tmin; k->tmax = tmax; k->qmin = qmin; k->qmax = qmax;
    k->time_only = time_only ? 1 : 0;
    TRYE(dev_alloc(&k->d_hitcount, nchannels));
    TRYE(dev_alloc(&k->d_tmom1, nchannels)); TRYE(dev_alloc(&k->d_tmom2, nchannels));
    TRYE(dev_alloc(&k->d_qmom1, nchannels)); TRYE(dev_alloc(&k->d_qmom2, nchannels));
    TRYE(dev_alloc(&k->d_event_hit, nchannels)); TRYE(dev_alloc(&k->d_event_time, nchannels));
    TRYE(dev_alloc(&k->d_event_charge, nchannels)); TRYE(dev_alloc(&k->d_inv_tbw, nchannels));
    TRYE(dev_alloc(&k->d_inv_qbw, nchannels)); TRYE(dev_alloc(&k->d_time_pdf, nchannels));
    TRYE(dev_alloc(&k->d_charge_pdf, nchannels));
    TRYE(chroma_kernel_pdf_clear_moments(k));
    *out = k;
    return 0;
}

extern "C" int chroma_kernel_pdf_accumulate_moments(chroma_kernel_pdf *k, const float *t_device,
                                                    const float *q_device) {
    ARG_CHECK(k && t_device && q_device);
    chroma_ctx *c = k->ctx;
    CTX_BIND(c);
    kpdf_moments_kernel<<<(k->nchannels + 255) / 256, 256, 0, c->stream>>>(
        k->time_only, (int)k->nchannels, t_device, q_device, k->tmin, k->tmax, k->qmin, k->qmax,
        k->d_hitcount, k->d_tmom1, k->d_tmom2, k->d_qmom1, k->d_qmom2);
    KERNEL_CHECK(c);
    return 0;
}

extern "C" int chroma_kernel_pdf_get_moments(chroma_kernel_pdf *k, uint32_t *hitcount, float *tmom1,
                                             float *tmom2, float *qmom1, float *qmom2) {
    ARG_CHECK(k);
    chroma_ctx *c = k->ctx;
    CTX_BIND(c);
    size_t nb = (size_t)k->nchannels * 4;
    if (hitcount) CU_CHECK(cudaMemcpyAsync(hitcount, k->d_hitcount, nb, cudaMemcpyDeviceToHost, c->stream));
    if (tmom1) CU_CHECK(cudaMemcpyAsync(tmom1, k->d_tmom1, nb, cudaMemcpyDeviceToHost, c->stream));
    if (tmom2) CU_CHECK(cudaMemcpyAsync(tmom2, k->d_tmom2, nb, cudaMemcpyDeviceToHost, c->stream));
    if (qmom1) CU_CHECK(cudaMemcpyAsync(qmom1, k->d_qmom1, nb, cudaMemcpyDeviceToHost, c->stream));
    if (qmom2) CU_CHECK(cudaMemcpyAsync(qmom2, k->d_qmom2, nb, cudaMemcpyDeviceToHost, c->stream));
    CU_CHECK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int chroma_kernel_pdf_clear_kernel(chroma_kernel_pdf *k) {
    ARG_CHECK(k);
    chroma_ctx *c = k->ctx;
    CTX_BIND(c);
    size_t nb = (size_t)k->nchannels * 4;
    CU_CHECK(cudaMemsetAsync(k->d_hitcount, 0, nb, c->stream));
    CU_CHECK(cudaMemsetAsync(k->d_time_pdf, 0, nb, c->stream));
    CU_CHECK(cudaMemsetAsync(k->d_charge_pdf, 0, nb, c->stream));
    return 0;
}

extern "C" int chroma_kernel_pdf_setup_kernel(chroma_kernel_pdf *k, const uint32_t *event_hit,
                                              const float *event_time, const float *event_charge,
                                              const float *inv_time_bandwidths,
                                              const float *inv_charge_bandwidths) {
    ARG_CHECK(k && event_hit && event_time && event_charge && inv_time_bandwidths);
    chroma_ctx *c = k->ctx;
    CTX_BIND(c);
    size_t nb = (size_t)k->nchannels * 4;
    CU_CHECK(cudaMemcpyAsync(k->d_event_hit, event_hit, nb, cudaMemcpyHostToDevice, c->stream));
    CU_CHECK(cudaMemcpyAsync(k->d_event_time, event_time, nb, cudaMemcpyHostToDevice, c->stream));
    CU_CHECK(cudaMemcpyAsync(k->d_event_charge, event_charge, nb, cudaMemcpyHostToDevice, c->stream));
    CU_CHECK(cudaMemcpyAsync(k->d_inv_tbw, inv_time_bandwidths, nb, cudaMemcpyHostToDevice, c->stream));
    if (inv_charge_bandwidths)
        CU_CHECK(cudaMemcpyAsync(k->d_inv_qbw, inv_charge_bandwidths, nb, cudaMemcpyHostToDevice, c->stream));
    else
        CU_CHECK(cudaMemsetAsync(k->d_inv_qbw, 0, nb, c->stream));
    int rc = chroma_kernel_pdf_clear_kernel(k);
    if (rc) return rc;
    CU_CHECK(cudaStreamSynchronize(c->stream));
    k->kernel_ready = true;
    return 0;
}

extern "C" int chroma_kernel_pdf_accumulate_kernel(chroma_kernel_pdf *k, const float *t_device,
                                                   const float *q_device) {
    ARG_CHECK(k && t_device && q_device);
    if (!k->kernel_ready) {
        chroma_set_error("setup_kernel() must be called before accumulate_kernel()");
        return CHROMA_ERR_STATE;
    }
    chroma_ctx *c = k->ctx;
    CTX_BIND(c);
    kpdf_eval_kernel<<<(k->nchannels + 255) / 256, 256, 0, c->stream>>>(
        k->time_only, (int)k->nchannels, k->d_event_hit, k->d_event_time, k->d_event_charge, t_device,
        q_device, k->tmin, k->tmax, k->qmin, k->qmax, k->d_inv_tbw, k->d_inv_qbw, k->d_hitcount,
        k->d_time_pdf, k->d_charge_pdf);
    KERNEL_CHECK(c);
    return 0;
}

extern "C" int chroma_kernel_pdf_get_kernel(chroma_kernel_pdf *k, uint32_t *hitcount, float *time_pdf_values,
                                            float *charge_pdf_values) {
    ARG_CHECK(k);
    chroma_ctx *c = k->ctx;
    CTX_BIND(c);
    size_t nb = (size_t)k->nchannels * 4;
    if (hitcount) CU_CHECK(cudaMemcpyAsync(hitcount, k->d_hitcount, nb, cudaMemcpyDeviceToHost, c->stream));
    if (time_pdf_values) CU_CHECK(cudaMemcpyAsync(time_pdf_values, k->d_time_pdf, nb, cudaMemcpyDeviceToHost, c->stream));
    if (charge_pdf_values) CU_CHECK(cudaMemcpyAsync(charge_pdf_values, k->d_charge_pdf, nb, cudaMemcpyDeviceToHost, c->stream));
    CU_CHECK(cudaStreamSynchronize(c->stream));
    return 0;
}
