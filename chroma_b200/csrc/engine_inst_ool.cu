// engine_inst_ool.cu — second build of the two-level propagate kernel with the random draws out of line
// (see wide_prop_kernel.cuh / inst_prop_kernel.cuh).
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
