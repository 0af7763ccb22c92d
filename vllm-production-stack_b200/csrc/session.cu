// session.cu -- batched session routing on sm_100a: MD5 of the session key,
// then bisect_right over the sorted 128-bit ring points, wrapping to point 0.
//
// Replaces uhashring.HashRing.get_node as called by SessionRouter.route_request
// (src/vllm_router/routers/routing_logic.py:217; ring maintenance :84-103).
// uhashring's default ring hashes with int(md5(key).hexdigest(), 16): the
// digest bytes read as one big-endian 128-bit integer.
//
// ALU-bound (64 MD5 steps per key, ~50 bytes of traffic), one thread per key;
// the ring (2560 points for 16 endpoints) is staged in shared memory.
#include "common.cuh"

namespace b200r {

__device__ __forceinline__ uint32_t rotl32(uint32_t x, int r) { return __funnelshift_l(x, x, r); }

__constant__ uint32_t MD5_K[64] = {
    0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501,
    0x698098d8, 0x8b44f7af, 0xffff5bb1, 0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821,
    0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa, 0xd62f105d, 0x02441453, 0xd8a1e681, 0xe7d3fbc8,
    0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8, 0x676f02d9, 0x8d2a4c8a,
    0xfffa3942, 0x8771f681, 0x6d9d6122, 0xfde5380c, 0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70,
    0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05, 0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665,
    0xf4292244, 0x432aff97, 0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d, 0x85845dd1,
    0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1, 0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391 };

// One 64-byte block; fully unrolled so m[] stays in registers.
__device__ __forceinline__ void md5_block(uint32_t st[4], const uint32_t m[16])
{
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3];
#define MD5_STEP(F, G, S, I)                                       \
    {                                                              \
        uint32_t x = a + (F) + MD5_K[I] + m[G];                    \
        a = d; d = c; c = b;                                       \
        b = b + rotl32(x, S);                                      \
    }
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const int s = (i & 3) == 0 ? 7 : (i & 3) == 1 ? 12 : (i & 3) == 2 ? 17 : 22;
        MD5_STEP((b & c) | (~b & d), i, s, i)
    }
#pragma unroll
    for (int i = 16; i < 32; i++) {
        const int s = (i & 3) == 0 ? 5 : (i & 3) == 1 ? 9 : (i & 3) == 2 ? 14 : 20;
        MD5_STEP((d & b) | (~d & c), (5 * i + 1) & 15, s, i)
    }
#pragma unroll
    for (int i = 32; i < 48; i++) {
        const int s = (i & 3) == 0 ? 4 : (i & 3) == 1 ? 11 : (i & 3) == 2 ? 16 : 23;
        MD5_STEP(b ^ c ^ d, (3 * i + 5) & 15, s, i)
    }
#pragma unroll
    for (int i = 48; i < 64; i++) {
        const int s = (i & 3) == 0 ? 6 : (i & 3) == 1 ? 10 : (i & 3) == 2 ? 15 : 21;
        MD5_STEP(c ^ (b | ~d), (7 * i) & 15, s, i)
    }
#undef MD5_STEP
    st[0] += a; st[1] += b; st[2] += c; st[3] += d;
}

// byte `o` of the padded message of a key of `len` bytes
__device__ __forceinline__ uint32_t msg_byte(const uint8_t *p, int len, int64_t padded, int64_t o)
{
    if (o < len) return p[o];
    if (o == len) return 0x80;
    if (o >= padded - 8) return (uint32_t)((((uint64_t)len * 8) >> (8 * (o - (padded - 8)))) & 0xFF);
    return 0;
}

// MD5 of p[0..len) as the big-endian 128-bit integer (hi, lo).
__device__ __forceinline__ void md5_u128(const uint8_t *p, int len, uint64_t &hi, uint64_t &lo)
{
    uint32_t st[4] = { 0x67452301u, 0xefcdab89u, 0x98badcfeu, 0x10325476u };
    const int64_t padded = ((int64_t)len + 8) / 64 * 64 + 64;
    const bool al4 = (((uintptr_t)p) & 3) == 0;
    for (int64_t blk = 0; blk < padded; blk += 64) {
        uint32_t m[16];
#pragma unroll
        for (int w = 0; w < 16; w++) {
            const int64_t o = blk + 4 * w;
            if (al4 && o + 4 <= len) {
                m[w] = *(const uint32_t *)(p + o);
            } else if (o >= len + 1 && o + 4 <= padded - 8) {
                m[w] = 0;
            } else {
                m[w] = msg_byte(p, len, padded, o) | (msg_byte(p, len, padded, o + 1) << 8) |
                       (msg_byte(p, len, padded, o + 2) << 16) | (msg_byte(p, len, padded, o + 3) << 24);
            }
        }
        md5_block(st, m);
    }
    // digest bytes = st[0..3] little-endian; read them as one big-endian integer
    hi = ((uint64_t)__byte_perm(st[0], 0, 0x0123) << 32) | __byte_perm(st[1], 0, 0x0123);
    lo = ((uint64_t)__byte_perm(st[2], 0, 0x0123) << 32) | __byte_perm(st[3], 0, 0x0123);
}

__global__ void k_md5_batch(const uint8_t *__restrict__ bytes, const int64_t *__restrict__ starts,
                            const int32_t *__restrict__ lens, int64_t n, uint64_t *__restrict__ out)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t hi, lo;
    md5_u128(bytes + starts[i], lens[i], hi, lo);
    out[2 * i] = hi;
    out[2 * i + 1] = lo;
}

// Ring points live in dynamic shared memory when they fit (hi[] then lo[]).
template <bool IN_SMEM>
__global__ void k_ring_lookup(const uint64_t *__restrict__ ring_hi, const uint64_t *__restrict__ ring_lo,
                              const int32_t *__restrict__ owner, int npoints,
                              const uint8_t *__restrict__ bytes, const int64_t *__restrict__ starts,
                              const int32_t *__restrict__ lens, int64_t n, int32_t *__restrict__ out)
{
    extern __shared__ __align__(16) uint64_t s_ring[];
    const uint64_t *rh = ring_hi, *rl = ring_lo;
    if (IN_SMEM) {
        for (int k = threadIdx.x; k < npoints; k += blockDim.x) {
            s_ring[k] = ring_hi[k];
            s_ring[npoints + k] = ring_lo[k];
        }
        __syncthreads();
        rh = s_ring;
        rl = s_ring + npoints;
    }
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        uint64_t hi, lo;
        md5_u128(bytes + starts[i], lens[i], hi, lo);
        int a = 0, b = npoints;                    // bisect_right
        while (a < b) {
            const int mid = (a + b) >> 1;
            const uint64_t mh = rh[mid];
            const bool less = (hi < mh) || (hi == mh && lo < rl[mid]);
            if (less) b = mid; else a = mid + 1;
        }
        if (a == npoints) a = 0;                   // wrap
        out[i] = owner[a];
    }
}

}  // namespace b200r

using namespace b200r;

struct b200r_ring {
    uint64_t *hi, *lo;
    int32_t *owner;
    int npoints, cap;
};

extern "C" {

int b200r_md5_batch(const uint8_t *d_bytes, const int64_t *d_starts, const int32_t *d_lens,
                    int64_t n, uint64_t *d_out, void *stream)
{
    if (n < 0) { set_error("b200r_md5_batch: n < 0"); return B200R_E_INVALID; }
    if (n == 0) return B200R_OK;
    if (!d_bytes || !d_starts || !d_lens || !d_out) { set_error("null pointer"); return B200R_E_INVALID; }
    const int blocks = (int)((n + 127) / 128);
    k_md5_batch<<<blocks, 128, 0, (cudaStream_t)stream>>>(d_bytes, d_starts, d_lens, n, d_out);
    B200R_LAUNCH_CHECK("k_md5_batch");
    return B200R_OK;
}

int b200r_ring_create(b200r_ring **out)
{
    if (!out) { set_error("null out"); return B200R_E_INVALID; }
    b200r_ring *r = new b200r_ring();
    r->hi = r->lo = nullptr; r->owner = nullptr; r->npoints = 0; r->cap = 0;
    *out = r;
    return B200R_OK;
}

int b200r_ring_destroy(b200r_ring *r)
{
    if (!r) return B200R_OK;
    cudaFree(r->hi); cudaFree(r->lo); cudaFree(r->owner);
    delete r;
    return B200R_OK;
}

int b200r_ring_build(b200r_ring *r, const uint64_t *h_hi, const uint64_t *h_lo,
                     const int32_t *h_owner, int32_t npoints, void *stream)
{
    if (!r || npoints < 0 || (npoints > 0 && (!h_hi || !h_lo || !h_owner))) {
        set_error("b200r_ring_build: bad argument");
        return B200R_E_INVALID;
    }
    for (int i = 1; i < npoints; i++) {
        if (h_hi[i - 1] > h_hi[i] || (h_hi[i - 1] == h_hi[i] && h_lo[i - 1] >= h_lo[i])) {
            set_error("b200r_ring_build: points must be strictly ascending (index %d)", i);
            return B200R_E_INVALID;
        }
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (npoints > r->cap) {
        B200R_CUDA(cudaStreamSynchronize(st));
        cudaFree(r->hi); cudaFree(r->lo); cudaFree(r->owner);
        r->hi = r->lo = nullptr; r->owner = nullptr; r->cap = 0;
        int cap = npoints < 4096 ? 4096 : npoints;
        B200R_CUDA(cudaMalloc(&r->hi, (size_t)cap * 8));
        B200R_CUDA(cudaMalloc(&r->lo, (size_t)cap * 8));
        B200R_CUDA(cudaMalloc(&r->owner, (size_t)cap * 4));
        r->cap = cap;
    }
    if (npoints > 0) {
        // pageable host memory: these copies return after the source is consumed
        B200R_CUDA(cudaMemcpyAsync(r->hi, h_hi, (size_t)npoints * 8, cudaMemcpyHostToDevice, st));
        B200R_CUDA(cudaMemcpyAsync(r->lo, h_lo, (size_t)npoints * 8, cudaMemcpyHostToDevice, st));
        B200R_CUDA(cudaMemcpyAsync(r->owner, h_owner, (size_t)npoints * 4, cudaMemcpyHostToDevice, st));
        B200R_CUDA(cudaStreamSynchronize(st));
    }
    r->npoints = npoints;
    return B200R_OK;
}

int b200r_ring_lookup_batch(b200r_ring *r, const uint8_t *d_bytes, const int64_t *d_starts,
                            const int32_t *d_lens, int64_t n, int32_t *d_endpoint, void *stream)
{
    if (!r || n < 0) { set_error("b200r_ring_lookup_batch: bad argument"); return B200R_E_INVALID; }
    if (n == 0) return B200R_OK;
    if (!d_bytes || !d_starts || !d_lens || !d_endpoint) { set_error("null pointer"); return B200R_E_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    if (r->npoints == 0) {          // HashRing.get_node on an empty ring returns None
        B200R_CUDA(cudaMemsetAsync(d_endpoint, 0xFF, (size_t)n * 4, st));
        return B200R_OK;
    }
    const int threads = 256;
    int64_t want = (n + threads - 1) / threads;
    const int blocks = (int)(want < 148 * 8 ? want : 148 * 8);
    const size_t smem = (size_t)r->npoints * 16;
    if (smem <= 200 * 1024) {
        static bool attr_set[64] = { false };     // the attribute is per device
        int devid = 0;
        B200R_CUDA(cudaGetDevice(&devid));
        if (devid >= 64 || !attr_set[devid]) {
            B200R_CUDA(cudaFuncSetAttribute(k_ring_lookup<true>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
            if (devid < 64) attr_set[devid] = true;
        }
        k_ring_lookup<true><<<blocks, threads, smem, st>>>(r->hi, r->lo, r->owner, r->npoints, d_bytes,
                                                          d_starts, d_lens, n, d_endpoint);
    } else {
        k_ring_lookup<false><<<blocks, threads, 0, st>>>(r->hi, r->lo, r->owner, r->npoints, d_bytes,
                                                        d_starts, d_lens, n, d_endpoint);
    }
    B200R_LAUNCH_CHECK("k_ring_lookup");
    return B200R_OK;
}

}  // extern "C"
