/* oracle/ref_cpu_render.cpp — TEST INFRASTRUCTURE ONLY.  The reference's chroma/cuda/render.cu
 * compiled for the host (its own translation unit, as under nvcc); see oracle/ref_cpu_wrapper.cpp
 * for how.  -> oracle/_ref/libchroma_ref_cpu_render.so */
#include "shim.h"

#include "render.cu"      // chroma/cuda/render.cu

#include "ref_cpu/geometry.h"

/* chroma/cuda/render.cu:35-181: one thread per ray; dx [n][alpha_depth], dxlen [n], color [n][alpha_depth]
 * persist between calls (keep_last_render, chroma/gpu/render.py:44-68) */
extern "C" void ref_cpu_render(const oracle_geometry *o, const uint32_t *colors, uint64_t n, float *pos, float *dir,
                               unsigned int alpha_depth, unsigned int *pixels, float *dx, unsigned int *dxlen,
                               float *color, unsigned int bg_color) {
    RefGeometry R(o);
    R.g.colors = (unsigned int *)colors;
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t id = 0; id < (int64_t)n; id++) {
        as_thread((uint64_t)id);
        render((int)n, (float3 *)pos, (float3 *)dir, &R.g, alpha_depth, pixels, dx, dxlen, (float4 *)color, bg_color);
    }
}
