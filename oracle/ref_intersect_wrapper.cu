// ref_intersect_wrapper.cu — TEST INFRASTRUCTURE ONLY (oracle/).
//
// The reference's benchmark kernel distance_to_mesh (chroma/cuda/mesh.h:123-151) calls
// intersect_mesh() but drops the triangle index and never takes last_hit_triangle.  This
// translation unit includes the reference's own headers WHERE THEY LIE (-I $(REF)/chroma/cuda;
// nothing is copied) and adds the thin kernels the parity tests need around the reference's
// device functions, in the same style as the reference's test/*.cu wrappers
// (test/rotate_test.cu, test/test_sample_cdf.cu).  Built by `make -C oracle ref` into
// oracle/_ref/intersect.cubin with the reference's own options (--use_fast_math,
// chroma/gpu/tools.py:12).
#include "mesh.h"

extern "C" {

// first-hit triangle index + distance; distance is left untouched on a miss exactly like
// distance_to_mesh (callers pre-fill it)
__global__ void intersect_with_index(int nthreads, float3 *_origin, float3 *_direction,
                                     int *_last_hit, Geometry *g, int *_triangle,
                                     float *_distance)
{
    __shared__ Geometry sg;

    if (threadIdx.x == 0)
        sg = *g;

    __syncthreads();

    int id = blockIdx.x * blockDim.x + threadIdx.x;

    if (id >= nthreads)
        return;

    g = &sg;

    float3 origin = _origin[id];
    float3 direction = _direction[id];
    direction /= norm(direction);

    int last = _last_hit ? _last_hit[id] : -1;
    float distance;
    int triangle_index = intersect_mesh(origin, direction, g, distance, last);

    _triangle[id] = triangle_index;
    if (triangle_index != -1)
        _distance[id] = distance;
}

} // extern "C"
