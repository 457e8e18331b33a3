/* oracle/ref_cpu/curand_kernel.h — TEST INFRASTRUCTURE ONLY.
 * Host stand-in for the device-only cuRAND header the reference includes (chroma/cuda/random.h:4),
 * found first on the include path when the reference's own sources are compiled for the CPU
 * (oracle/ref_cpu_wrapper.cpp).  `curandState` is a Philox4x32-10 stream with the engine's keying
 * (seed; photon id, call epoch, stream), so the reference's code consumes the same uniforms, in its
 * own draw order, as the engine and the oracle port do. */
#ifndef CHROMA_REF_CPU_CURAND_SHIM_H
#define CHROMA_REF_CPU_CURAND_SHIM_H
#include <stdint.h>
#include <math.h>

struct curandState {
    uint32_t k0, k1, c0, c2, c3, blk, r[4], left;
};

static inline void refcpu_rng_seed(curandState *s, uint64_t seed, uint32_t id, uint32_t epoch, uint32_t stream) {
    s->k0 = (uint32_t)seed; s->k1 = (uint32_t)(seed >> 32);
    s->c0 = id; s->c2 = epoch; s->c3 = stream; s->blk = 0; s->left = 0;
}

static inline uint32_t refcpu_rng_u32(curandState *s) {
    if (s->left == 0) {
        uint32_t c0 = s->c0, c1 = s->blk, c2 = s->c2, c3 = s->c3, k0 = s->k0, k1 = s->k1;
        for (int i = 0; i < 10; i++) {
            const uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
            const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
            const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
            c0 = n0; c1 = n1; c2 = n2; c3 = n3;
            k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
        }
        s->r[0] = c0; s->r[1] = c1; s->r[2] = c2; s->r[3] = c3;
        s->blk++; s->left = 4;
    }
    return s->r[4 - s->left--];
}

/* curand_uniform: (0, 1] */
static inline float curand_uniform(curandState *s) {
    return fmaf((float)refcpu_rng_u32(s), 2.3283064365386963e-10f, 1.1641532182693481e-10f);
}

/* curand_normal: Box-Muller on two uniforms (sqrt(-2 ln u1) sin(2 pi u2)), the form the engine uses */
static inline float curand_normal(curandState *s) {
    const float u1 = curand_uniform(s), u2 = curand_uniform(s);
    return sqrtf(-2.0f * logf(u1)) * sinf(2.0f * 3.14159265358979323846f * u2);
}

/* chroma/cuda/random.h:60-70 init_rng: (seed, sequence = thread id, offset) */
static inline void curand_init(unsigned long long seed, unsigned long long sequence, unsigned long long offset,
                               curandState *s) {
    refcpu_rng_seed(s, seed, (uint32_t)sequence, (uint32_t)offset, 0u);
}
#endif
