// common.cuh -- shared device helpers for libb200router (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "b200_router.h"

namespace b200r {

// ---------------------------------------------------------------- host errors
void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what);

#define B200R_CUDA(call)                                       \
    do {                                                       \
        cudaError_t e__ = (call);                              \
        if (e__ != cudaSuccess) return cuda_fail(e__, #call);  \
    } while (0)

#define B200R_LAUNCH_CHECK(name)                               \
    do {                                                       \
        cudaError_t e__ = cudaGetLastError();                  \
        if (e__ != cudaSuccess) return cuda_fail(e__, name);   \
    } while (0)

// ---------------------------------------------------------------------- XXH64
// Bit-exact XXH64 (xxHash 0.8.x), seed 0: the hash the reference takes of every
// chunk (hashtrie.py:56).
constexpr uint64_t XP1 = 0x9E3779B185EBCA87ULL;
constexpr uint64_t XP2 = 0xC2B2AE3D27D4EB4FULL;
constexpr uint64_t XP3 = 0x165667B19E3779F9ULL;
constexpr uint64_t XP4 = 0x85EBCA77C2B2AE63ULL;
constexpr uint64_t XP5 = 0x27D4EB2F165667C5ULL;

// 64-bit helpers written against the 32-bit integer pipes of sm_100a: a 64x64->64
// multiply by a constant is IMAD.WIDE + 2 IMAD, a 64-bit rotate is two funnel shifts.
__device__ __forceinline__ void split64(uint64_t x, uint32_t &lo, uint32_t &hi)
{
    asm("mov.b64 {%0,%1}, %2;" : "=r"(lo), "=r"(hi) : "l"(x));
}
__device__ __forceinline__ uint64_t join64(uint32_t lo, uint32_t hi)
{
    uint64_t x;
    asm("mov.b64 %0, {%1,%2};" : "=l"(x) : "r"(lo), "r"(hi));
    return x;
}
template <int R>
__device__ __forceinline__ uint64_t rotl(uint64_t x)
{
    uint32_t lo, hi;
    split64(x, lo, hi);
    if (R >= 32) { uint32_t t = lo; lo = hi; hi = t; }
    constexpr int r = R & 31;
    if (r == 0) return join64(lo, hi);
    return join64(__funnelshift_l(hi, lo, r), __funnelshift_l(lo, hi, r));
}
__device__ __forceinline__ uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }
template <uint64_t C>
__device__ __forceinline__ uint64_t mulc(uint64_t x)
{
    uint32_t xl, xh, wl, wh;
    split64(x, xl, xh);
    uint64_t w;
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(w) : "r"(xl), "r"((uint32_t)C));
    split64(w, wl, wh);
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(wh) : "r"(xh), "r"((uint32_t)C));
    asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(wh) : "r"(xl), "r"((uint32_t)(C >> 32)));
    return join64(wl, wh);
}
__device__ __forceinline__ uint64_t xround(uint64_t acc, uint64_t in)
{
    return mulc<XP1>(rotl<31>(acc + mulc<XP2>(in)));
}
__device__ __forceinline__ uint64_t xmerge(uint64_t h, uint64_t v)
{
    return mulc<XP1>(h ^ xround(0, v)) + XP4;
}
__device__ __forceinline__ uint64_t xavalanche(uint64_t h)
{
    h ^= h >> 33; h = mulc<XP2>(h); h ^= h >> 29; h = mulc<XP3>(h); h ^= h >> 32;
    return h;
}
__device__ __forceinline__ uint64_t u64_of(uint32_t lo, uint32_t hi)
{
    return (uint64_t)lo | ((uint64_t)hi << 32);
}

// Unaligned-safe little-endian loads from global memory.  Only aligned words
// that contain at least one requested byte are touched, so nothing is read
// outside the allocation granule of the buffer.
__device__ __forceinline__ uint64_t ldg64_any(const uint8_t *p)
{
    uintptr_t a = (uintptr_t)p & 7;
    const uint64_t *q = (const uint64_t *)(p - a);
    uint64_t lo = q[0];
    if (a == 0) return lo;
    uint64_t hi = q[1];
    return (lo >> (8 * a)) | (hi << (64 - 8 * a));
}
__device__ __forceinline__ uint32_t ldg32_any(const uint8_t *p)
{
    return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}

// Generic XXH64 over global memory, any alignment, any length (slow path).
__device__ inline uint64_t xxh64_global(const uint8_t *p, int len)
{
    const uint8_t *end = p + len;
    uint64_t h;
    if (len >= 32) {
        uint64_t v1 = XP1 + XP2, v2 = XP2, v3 = 0, v4 = 0 - XP1;
        const uint8_t *lim = end - 32;
        do {
            v1 = xround(v1, ldg64_any(p));
            v2 = xround(v2, ldg64_any(p + 8));
            v3 = xround(v3, ldg64_any(p + 16));
            v4 = xround(v4, ldg64_any(p + 24));
            p += 32;
        } while (p <= lim);
        h = rotl<1>(v1) + rotl<7>(v2) + rotl<12>(v3) + rotl<18>(v4);
        h = xmerge(h, v1); h = xmerge(h, v2); h = xmerge(h, v3); h = xmerge(h, v4);
    } else {
        h = XP5;
    }
    h += (uint64_t)len;
    while (p + 8 <= end) { h ^= xround(0, ldg64_any(p)); h = mulc<XP1>(rotl<27>(h)) + XP4; p += 8; }
    if (p + 4 <= end) { h ^= mulc<XP1>((uint64_t)ldg32_any(p)); h = mulc<XP2>(rotl<23>(h)) + XP3; p += 4; }
    while (p < end) { h ^= mulc<XP5>((uint64_t)(*p)); h = mulc<XP1>(rotl<11>(h)); p++; }
    return xavalanche(h);
}

// --------------------------------------------------------------- tie-break
__device__ __host__ __forceinline__ uint32_t pick(uint64_t seed, uint64_t seq, uint32_t n)
{
    uint64_t x = seed + seq * 0x9E3779B97F4A7C15ULL;
    x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ULL;
    x ^= x >> 27; x *= 0x94D049BB133111EBULL;
    x ^= x >> 31;
    return (uint32_t)(((x >> 32) * (uint64_t)n) >> 32);
}

__device__ __forceinline__ int nth_set_bit(uint64_t m, uint32_t k)
{
    for (uint32_t i = 0; i < k; i++) m &= m - 1;
    return __ffsll((long long)m) - 1;
}

}  // namespace b200r
