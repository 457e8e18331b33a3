// engine_inst_ool.cu — second build of the two-level propagate kernel with the random draws out of line
// (see wide_prop_kernel.cuh / inst_prop_kernel.cuh).
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
#include "inst.cuh"
#define IPROP_KERNEL_NAME inst_propagate_ool_kernel
#include "inst_prop_kernel.cuh"

int chroma_inst_ool_occupancy(chroma_ctx *c, size_t smem, int *occ) {
    if (!c->inst_ool_attr_set) {
        CU_CHECK(cudaFuncSetAttribute(inst_propagate_ool_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU_CHECK(cudaFuncSetAttribute(inst_propagate_ool_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        c->inst_ool_attr_set = true;
    }
    CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, inst_propagate_ool_kernel<false>, IPROP_THREADS, smem));
    if (*occ < 1) *occ = 1;
    return 0;
}

void chroma_inst_ool_launch(chroma_geometry *g, chroma_photons *p, const WPropParams &prm, bool staged, unsigned grid,
                            size_t smem) {
    chroma_ctx *c = g->ctx;
    if (staged) inst_propagate_ool_kernel<true><<<grid, IPROP_THREADS, smem, c->stream>>>(g->wview, g->iv, g->wide, p->v, prm);
    else inst_propagate_ool_kernel<false><<<grid, IPROP_THREADS, smem, c->stream>>>(g->wview, g->iv, g->wide, p->v, prm);
}
