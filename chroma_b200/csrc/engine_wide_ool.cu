// engine_wide_ool.cu — second build of the wide-tree propagate kernel with the random draws out of line
// (see wide_prop_kernel.cuh); selected by chroma_wide_propagate for physics-heavy geometries.
// Search variants of this translation unit (r2f A/B on one B200, ms per 1e7-photon C3 bomb / per
// 2.5e6-photon C4 event; (two-pass group step, leaf spread) = 11 / 00 / 01 / 10):
//   flat C3 26.37 / 27.29 / 26.65 / 27.56    two-level C3 25.38 / 24.36 / 25.22 / 24.76
//   flat C4 19.94 / 18.33 / 19.21 / 18.86    two-level C4 23.40 / 22.05 / 24.29 / 22.63
// -> 11 only where the search of a huge flat tree dominates (engine_wide.cu); 00 here.
#ifndef WIDE_TWO_PASS
#define WIDE_TWO_PASS 0
#endif
#ifndef WIDE_LEAF_SPREAD
#define WIDE_LEAF_SPREAD 0
#endif
#define RNG_DRAW_NOINLINE 1
#include "physics.cuh"
#include "wide.cuh"
#define WPROP_KERNEL_NAME wide_propagate_ool_kernel
#include "wide_prop_kernel.cuh"

int chroma_wide_ool_occupancy(chroma_ctx *c, size_t smem, int *occ) {
    if (!c->wide_ool_attr_set) {
        CU_CHECK(cudaFuncSetAttribute(wide_propagate_ool_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU_CHECK(cudaFuncSetAttribute(wide_propagate_ool_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        c->wide_ool_attr_set = true;
    }
    CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, wide_propagate_ool_kernel<false>, WPROP_THREADS, smem));
    if (*occ < 1) *occ = 1;
    return 0;
}

void chroma_wide_ool_launch(chroma_geometry *g, chroma_photons *p, const WPropParams &prm, bool staged,
                            unsigned grid, size_t smem) {
    chroma_ctx *c = g->ctx;
    if (staged) wide_propagate_ool_kernel<true><<<grid, WPROP_THREADS, smem, c->stream>>>(g->wview, g->wide, p->v, prm);
    else wide_propagate_ool_kernel<false><<<grid, WPROP_THREADS, smem, c->stream>>>(g->wview, g->wide, p->v, prm);
}
