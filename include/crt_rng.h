/* crt_rng.h — counter-based per-(pixel, sample) random numbers.
 *
 * Replaces the reference's per-pixel XORWOW curandState
 * (CudaTest/src/core/deviceManage.h:120-136, 48 B of state per pixel) with a
 * stateless generator: draw k of the path that belongs to camera sample
 * `sample` of pixel `pixel` is a pure function of (seed, pixel, sample, k).
 * The only thing a path carries is the 32-bit draw counter k.
 *
 * Like curand_uniform (used at render.h:116-117, camera.h:10,
 * material.h:14-16,134) the float result lies in (0, 1].
 *
 * Integer-only arithmetic, so the CUDA kernels, the C oracle and the shim
 * that feeds the unmodified reference headers all produce the same stream.
 * Plain C; also valid C++ and CUDA (define CRT_HD before including).
 */
#ifndef CRT_RNG_H
#define CRT_RNG_H

#include <stdint.h>

#ifndef CRT_HD
#  ifdef __CUDACC__
#    define CRT_HD __host__ __device__ __forceinline__
#  else
#    define CRT_HD static inline
#  endif
#endif

/* 32-bit finaliser ("lowbias32" constants); a bijection on uint32. */
CRT_HD uint32_t crt_mix32(uint32_t x)
{
    x ^= x >> 16;
    x *= 0x7feb352dU;
    x ^= x >> 15;
    x *= 0x846ca68bU;
    x ^= x >> 16;
    return x;
}

/* Stream keys are 64 bits wide, held as one uint64 = (k1 << 32) | k0, and the map
 * (seed, pixel, sample) -> key is INJECTIVE for a fixed seed: k0 = mix(pixel ^ a) and
 * k1 - k0 = mix(sample ^ b) are bijections of the pixel and of the sample number, so no two camera samples of a
 * frame share a stream, whatever the frame size (a 32-bit key made whole sample planes of a 4K x 64 spp frame
 * shifted copies of each other). (a, b) is a bijection of the 64-bit seed, so two seeds never give the same
 * frame either.
 */

/* The seed's share of a stream key: the same for every sample of a frame, so a renderer computes it once. */
CRT_HD uint64_t crt_rng_seed_key(uint64_t seed)
{
    const uint32_t a = crt_mix32((uint32_t)seed ^ 0x9e3779b9U);
    const uint32_t b = crt_mix32((uint32_t)(seed >> 32) + a);
    return ((uint64_t)b << 32) | a;
}

/* Stream key of one camera sample from the seed's share. */
CRT_HD uint64_t crt_rng_key_from(uint64_t seed_key, uint32_t pixel, uint32_t sample)
{
    const uint32_t k0 = crt_mix32(pixel ^ (uint32_t)seed_key);
    const uint32_t k1 = crt_mix32(sample ^ (uint32_t)(seed_key >> 32)) + k0;
    return ((uint64_t)k1 << 32) | k0;
}

/* Stream key of one camera sample. */
CRT_HD uint64_t crt_rng_key(uint64_t seed, uint32_t pixel, uint32_t sample)
{
    return crt_rng_key_from(crt_rng_seed_key(seed), pixel, sample);
}

/* Raw 32-bit draw number `ctr` of stream `key`: both key words enter, one per mixing round. */
CRT_HD uint32_t crt_rng_bits(uint64_t key, uint32_t ctr)
{
    return crt_mix32((uint32_t)key ^ crt_mix32((uint32_t)(key >> 32) ^ (ctr * 0x9e3779b1U + 0x7f4a7c15U)));
}

/* Uniform float in (0, 1]: 24 random bits, never 0, may be exactly 1. */
CRT_HD float crt_rng_uniform(uint64_t key, uint32_t ctr)
{
    return (float)((crt_rng_bits(key, ctr) >> 8) + 1U) * (1.0f / 16777216.0f);
}

#endif /* CRT_RNG_H */
