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

// ------------------------------------------------------------------ path keys
// A trie node is identified by the sequence of chunk hashes from the root
// (hashtrie.py:70-72).  The flat table keys a node by a 128-bit polynomial
// path key, one 64-bit lane per multiplier:
//     K1_d = A1 * K1_{d-1} + h_d              K1_0 = ROOT1
//     K2_d = A2 * K2_{d-1} + rotl(h_d, 32)    K2_0 = ROOT2       (mod 2^64)
// Affine in the parent key, so (a) a warp computes all depths of a request with
// a 5-step Kogge-Stone scan (constants A^(2^k)), and (b) the recurrence is
// invertible given the stored chunk hash (A1, A2 odd), which lets a dump be
// decoded back into the pointer trie.
constexpr uint64_t KA1 = 0x9E3779B97F4A7C15ULL;
constexpr uint64_t KA2 = 0xD6E8FEB86659FD93ULL;
constexpr uint64_t KROOT1 = 0x243F6A8885A308D3ULL;
constexpr uint64_t KROOT2 = 0x13198A2E03707344ULL;

__host__ __device__ constexpr uint64_t cpow(uint64_t a, int e)
{
    uint64_t r = 1;
    for (int i = 0; i < e; i++) r *= a;
    return r;
}
__host__ __device__ constexpr uint64_t cinv(uint64_t a)
{   // Newton iteration for the inverse of an odd number mod 2^64
    uint64_t x = a;
    for (int i = 0; i < 6; i++) x *= 2 - a * x;
    return x;
}

struct Key { uint64_t k1, k2; };

// Inclusive scan of the key recurrence over the 32 lanes of a warp.
// In: h = this lane's chunk hash (0 for lanes past the end), carry = key of the
// node just before lane 0.  Out: key of this lane's node.
__device__ __forceinline__ Key key_scan_warp(uint64_t h, Key carry, int lane)
{
    uint64_t p1 = h, p2 = rotl<32>(h);
    if (lane == 0) { p1 += KA1 * carry.k1; p2 += KA2 * carry.k2; }
#define B200R_KS_STEP(S)                                           \
    {                                                              \
        uint64_t t1 = __shfl_up_sync(0xffffffffu, p1, S);          \
        uint64_t t2 = __shfl_up_sync(0xffffffffu, p2, S);          \
        constexpr uint64_t m1 = cpow(KA1, S), m2 = cpow(KA2, S);   \
        if (lane >= S) { p1 += m1 * t1; p2 += m2 * t2; }           \
    }
    B200R_KS_STEP(1) B200R_KS_STEP(2) B200R_KS_STEP(4) B200R_KS_STEP(8) B200R_KS_STEP(16)
#undef B200R_KS_STEP
    Key k; k.k1 = p1; k.k2 = p2;
    return k;
}

// -------------------------------------------------------------- table entries
// One entry = one 64-byte DRAM access atom.  The first 32-byte sector holds everything a
// probe needs (key, endpoint mask, batch bookkeeping); the second sector holds the chunk
// hash of the edge parent -> node, which is written once at creation and read only by
// dumps.  L2 fetches HBM in 64-byte atoms, so after the probe's miss that second sector is
// already resident and the store is an L2 hit -- a separate side array cost one more random
// DRAM read-modify-write per new node (measured: claim 51 -> 45 us per 64K batch).
struct __align__(64) Entry {
    uint64_t k1, k2;   // path key; (0,0) = empty.  Stored keys have k1,k2 != 0.
    uint64_t mask;     // endpoint set
    unsigned long long first;   // batch bookkeeping: (epoch << 32) | ~request of the first
                                // request of batch `epoch` that touched the node (atomicMax)
    uint64_t h;        // chunk hash of the edge parent -> node (dumps / path reconstruction)
    uint64_t reserved[3];
};

struct TrieDev {             // device-resident control block
    uint64_t root_mask;
    unsigned long long num_nodes;    // filled by k_count_nodes (status / dump), not on the hot path
    unsigned int status;
    unsigned int dump_count_lo;
    unsigned long long dump_count;
};

__device__ __forceinline__ uint32_t slot_of(uint64_t k1, uint64_t k2, int cap_log2)
{
    uint64_t x = (k1 ^ (k2 >> 32)) * 0x9E3779B97F4A7C15ULL;
    return (uint32_t)(x >> (64 - cap_log2));
}

__device__ __forceinline__ void ld_key(const Entry *e, uint64_t &a, uint64_t &b)
{   // L2-coherent 16-byte read
    asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(e) : "memory");
}
__device__ __forceinline__ void ld_key_l1(const Entry *e, uint64_t &a, uint64_t &b)
{   // L1-cached 16-byte read of a key.  Keys never change once written, so a stale
    // line can only show "empty" where a key now lives -- and an empty slot is always
    // confirmed by the CAS, which returns the truth.  Hot nodes (shared prefixes, probed
    // by thousands of requests per batch) are then served from L1 instead of L2.
    asm volatile("ld.global.ca.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(e) : "memory");
}

__device__ __forceinline__ void cas128(Entry *e, uint64_t n1, uint64_t n2, uint64_t &o1, uint64_t &o2)
{   // compare with (0,0), swap in (n1,n2); returns the previous key
    asm volatile(
        "{\n\t"
        ".reg .b128 cmp, nw, old;\n\t"
        "mov.b128 cmp, {%2, %3};\n\t"
        "mov.b128 nw, {%4, %5};\n\t"
        "atom.global.relaxed.gpu.cas.b128 old, [%6], cmp, nw;\n\t"
        "mov.b128 {%0, %1}, old;\n\t"
        "}\n"
        : "=l"(o1), "=l"(o2)
        : "l"(0ULL), "l"(0ULL), "l"(n1), "l"(n2), "l"(e)
        : "memory");
}

constexpr uint32_t SLOT_NONE = 0xFFFFFFFFu;

// Probe for key (k1,k2).  Returns the slot or SLOT_NONE (absent).
__device__ __forceinline__ uint32_t table_find(const Entry *tab, int cap_log2, uint64_t k1, uint64_t k2)
{
    if (k1 == 0) k1 = 1;
    if (k2 == 0) k2 = 1;
    const uint32_t m = (1u << cap_log2) - 1;
    uint32_t s = slot_of(k1, k2, cap_log2);
    for (uint32_t probe = 0; probe <= m; probe++) {
        uint64_t a, b;
        ld_key_l1(tab + s, a, b);
        if (a == k1 && b == k2) return s;
        if (a == 0 && b == 0) return SLOT_NONE;
        s = (s + 1) & m;
    }
    return SLOT_NONE;
}

// Find or create the entry for (k1,k2).  Returns the slot, SLOT_NONE if the
// table is full.  *created is set when this call made the entry.
__device__ __forceinline__ uint32_t table_claim(Entry *tab, uint64_t *hs, int cap_log2, uint64_t k1,
                                                uint64_t k2, uint64_t h, bool *created)
{
    if (k1 == 0) k1 = 1;
    if (k2 == 0) k2 = 1;
    const uint32_t m = (1u << cap_log2) - 1;
    uint32_t s = slot_of(k1, k2, cap_log2);
    *created = false;
    for (uint32_t probe = 0; probe <= m; probe++) {
        Entry *e = tab + s;
        uint64_t a, b;
        ld_key_l1(e, a, b);
        for (;;) {
            if (a == k1 && b == k2) return s;
            if (a == 0 && b == 0) {
                cas128(e, k1, k2, a, b);
                if (a == 0 && b == 0) {
                    if (hs) e->h = h;
                    *created = true;
                    return s;
                }
                continue;           // somebody else took it: re-examine (a,b)
            }
            if ((a == k1 && b == 0) || (a == 0 && b == k2)) {   // torn read: retry
                ld_key(e, a, b);
                continue;
            }
            break;                  // other key
        }
        s = (s + 1) & m;
    }
    return SLOT_NONE;
}

// One 32-byte load of the probe sector {k1, k2, mask, first} (LDG.256, sm_100): a probe
// that finds its node needs no second access for the mask and the first-toucher word.
__device__ __forceinline__ void ld_entry_l1(const Entry *e, uint64_t &a, uint64_t &b, uint64_t &mask,
                                            uint64_t &first)
{
    asm volatile("ld.global.ca.v4.u64 {%0, %1, %2, %3}, [%4];"
                 : "=l"(a), "=l"(b), "=l"(mask), "=l"(first) : "l"(e) : "memory");
}

// table_claim that also returns the node's mask and first-toucher word as seen by the
// probe.  For a node that did not exist when it was probed both are 0, which is also the
// truth for the rest of the claim kernel (masks only change in the commit kernel; a stale
// `first` is only a hint that saves an atomicMax).
__device__ __forceinline__ uint32_t table_claim_mf(Entry *tab, int cap_log2, uint64_t k1, uint64_t k2,
                                                   uint64_t h, bool *created, uint64_t *mask,
                                                   uint64_t *first)
{
    if (k1 == 0) k1 = 1;
    if (k2 == 0) k2 = 1;
    const uint32_t m = (1u << cap_log2) - 1;
    uint32_t s = slot_of(k1, k2, cap_log2);
    *created = false;
    for (uint32_t probe = 0; probe <= m; probe++) {
        Entry *e = tab + s;
        uint64_t a, b;
        ld_entry_l1(e, a, b, *mask, *first);
        for (;;) {
            if (a == k1 && b == k2) return s;
            if (a == 0 && b == 0) {
                cas128(e, k1, k2, a, b);
                *mask = 0; *first = 0;
                if (a == 0 && b == 0) {
                    e->h = h;
                    *created = true;
                    return s;
                }
                continue;           // somebody else took it: re-examine (a,b)
            }
            if ((a == k1 && b == 0) || (a == 0 && b == k2)) {   // torn read: retry
                ld_key(e, a, b);
                *mask = 0; *first = 0;
                continue;
            }
            break;                  // other key
        }
        s = (s + 1) & m;
    }
    return SLOT_NONE;
}

// A^(n+1) for n = 0..31, both key lanes: carry multipliers of a 32-chunk tile
#define B200R_KP(N) { cpow(KA1, N), cpow(KA2, N) }
__device__ const ulonglong2 g_key_pow[32] = {
    B200R_KP(1), B200R_KP(2), B200R_KP(3), B200R_KP(4), B200R_KP(5), B200R_KP(6), B200R_KP(7), B200R_KP(8),
    B200R_KP(9), B200R_KP(10), B200R_KP(11), B200R_KP(12), B200R_KP(13), B200R_KP(14), B200R_KP(15), B200R_KP(16),
    B200R_KP(17), B200R_KP(18), B200R_KP(19), B200R_KP(20), B200R_KP(21), B200R_KP(22), B200R_KP(23), B200R_KP(24),
    B200R_KP(25), B200R_KP(26), B200R_KP(27), B200R_KP(28), B200R_KP(29), B200R_KP(30), B200R_KP(31), B200R_KP(32) };
#undef B200R_KP

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
