/* oracle/ref_cpu_pdf.cpp — TEST INFRASTRUCTURE ONLY.  The reference's chroma/cuda/pdf.cu compiled
 * for the host (its own translation unit, as under nvcc: sorting.h has no include guard); see
 * oracle/ref_cpu_wrapper.cpp for how.  -> oracle/_ref/libchroma_ref_cpu_pdf.so */
#include "shim.h"

#include "pdf.cu"         // chroma/cuda/pdf.cu

/* chroma/cuda/pdf.cu:9-32 bin_hits, one thread per channel (chroma/gpu/pdf.py:211-227) */
extern "C" void ref_cpu_bin_hits(int nchannels, float *channel_q, float *channel_time, unsigned int *hitcount,
                                 int tbins, float tmin, float tmax, int qbins, float qmin, float qmax,
                                 unsigned int *pdf) {
    for (int id = 0; id < nchannels; id++) {
        as_thread(id);
        bin_hits(nchannels, channel_q, channel_time, hitcount, tbins, tmin, tmax, qbins, qmin, qmax, pdf);
    }
}

/* chroma/cuda/pdf.cu:34-219 under the host sequence of chroma/gpu/pdf.py:229-327: the channel <-> hit
 * maps of setup_pdf_eval, work queues filled with 1, accumulate_bincount over the channels, then
 * accumulate_nearest_neighbor_block with one block per hit channel */
extern "C" void ref_cpu_accumulate_pdf_eval(int nchannels, int ndaq, unsigned int *event_hit, float *event_time,
                                            float *mc_time, unsigned int *hitcount, unsigned int *bincount,
                                            float min_twidth, float tmin, float tmax, int min_bin_content,
                                            float *nearest_mc) {
    std::vector<unsigned int> ch_to_hit(nchannels, 0u), hit_to_ch;
    for (int c = 0; c < nchannels; c++)
        if (event_hit[c]) { ch_to_hit[c] = (unsigned)hit_to_ch.size(); hit_to_ch.push_back((unsigned)c); }
    const int nhit = (int)hit_to_ch.size();
    std::vector<unsigned int> work((size_t)(nhit > 0 ? nhit : 1) * (ndaq + 1), 1u);
    for (int c = 0; c < nchannels; c++) {
        as_thread(c);
        accumulate_bincount(nchannels, ndaq, event_hit, event_time, mc_time, hitcount, bincount, min_twidth, tmin,
                            tmax, min_bin_content, ch_to_hit.data(), work.data());
    }
    for (int h = 0; h < nhit; h++) {
        as_thread(h);
        accumulate_nearest_neighbor_block(nhit, ndaq, hit_to_ch.data(), work.data(), event_time, mc_time,
                                          nearest_mc, min_bin_content);
    }
}

/* chroma/cuda/pdf.cu:223-266 */
extern "C" void ref_cpu_accumulate_moments(int time_only, int nchannels, float *mc_time, float *mc_charge,
                                           float tmin, float tmax, float qmin, float qmax, unsigned int *mom0,
                                           float *t_mom1, float *t_mom2, float *q_mom1, float *q_mom2) {
    for (int id = 0; id < nchannels; id++) {
        as_thread(id);
        accumulate_moments(time_only, nchannels, mc_time, mc_charge, tmin, tmax, qmin, qmax, mom0, t_mom1, t_mom2,
                           q_mom1, q_mom2);
    }
}

/* chroma/cuda/pdf.cu:268-368 */
extern "C" void ref_cpu_accumulate_kernel_eval(int time_only, int nchannels, unsigned int *event_hit,
                                               float *event_time, float *event_charge, float *mc_time,
                                               float *mc_charge, float tmin, float tmax, float qmin, float qmax,
                                               float *inv_tbw, float *inv_qbw, unsigned int *hitcount,
                                               float *time_pdf, float *charge_pdf) {
    for (int id = 0; id < nchannels; id++) {
        as_thread(id);
        accumulate_kernel_eval(time_only, nchannels, event_hit, event_time, event_charge, mc_time, mc_charge, tmin,
                               tmax, qmin, qmax, inv_tbw, inv_qbw, hitcount, time_pdf, charge_pdf);
    }
}

