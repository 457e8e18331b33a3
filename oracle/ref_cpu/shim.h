/* oracle/ref_cpu/shim.h — TEST INFRASTRUCTURE ONLY.
 * What lets the reference's CUDA sources compile as host C++ (see oracle/ref_cpu_wrapper.cpp):
 * CUDA's host-compilable vector / complex headers, the execution-space qualifiers defined away,
 * thread-local threadIdx/blockIdx/blockDim, __syncthreads() as a no-op (kernels are called one
 * "thread" per block), atomics as GCC builtins, min/max overloads of the CUDA math API. */
#ifndef CHROMA_REF_CPU_SHIM_H
#define CHROMA_REF_CPU_SHIM_H
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <cmath>
#include <vector>
#include <omp.h>

#include <vector_types.h>
#include <vector_functions.h>
#include <cuComplex.h>

#undef __device__
#undef __global__
#undef __host__
#undef __shared__
#undef __noinline__
#define __device__
#define __global__
#define __host__
#define __shared__
#define __noinline__

using std::isnan;
using std::isfinite;

static thread_local uint3 threadIdx = {0, 0, 0}, blockIdx = {0, 0, 0};
static thread_local uint3 blockDim = {1, 1, 1}, gridDim = {1, 1, 1};
static inline void __syncthreads() {}
static inline float __int_as_float(int i) { float f; memcpy(&f, &i, 4); return f; }
static inline int __float_as_int(float f) { int i; memcpy(&i, &f, 4); return i; }
static inline float max(float a, float b) { return fmaxf(a, b); }
static inline float min(float a, float b) { return fminf(a, b); }
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }
static inline unsigned int min(unsigned int a, unsigned int b) { return a < b ? a : b; }
static inline unsigned int max(unsigned int a, unsigned int b) { return a > b ? a : b; }
static inline unsigned int atomicAdd(unsigned int *p, unsigned int v) { return __atomic_fetch_add(p, v, __ATOMIC_RELAXED); }
static inline int atomicAdd(int *p, int v) { return __atomic_fetch_add(p, v, __ATOMIC_RELAXED); }
static inline unsigned int atomicOr(unsigned int *p, unsigned int v) { return __atomic_fetch_or(p, v, __ATOMIC_RELAXED); }
static inline int atomicMin(int *p, int v) {
    int old = __atomic_load_n(p, __ATOMIC_RELAXED);
    while (v < old && !__atomic_compare_exchange_n(p, &old, v, true, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {}
    return old;
}
static inline unsigned int atomicMin(unsigned int *p, unsigned int v) {
    unsigned int old = __atomic_load_n(p, __ATOMIC_RELAXED);
    while (v < old && !__atomic_compare_exchange_n(p, &old, v, true, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {}
    return old;
}


static inline void as_thread(uint64_t id) {
    blockIdx.x = (unsigned)id; blockDim.x = 1; threadIdx.x = 0;
}
#endif
