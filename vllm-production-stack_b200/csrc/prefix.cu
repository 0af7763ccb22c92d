// prefix.cu -- batched prefix-aware routing on sm_100a.
//
// Replaces, behind the C ABI of include/b200_router.h, the reference's
//   HashTrie._chunk_and_hash / longest_prefix_match / insert
//       (src/vllm_router/prefix/hashtrie.py:45-103)
//   PrefixAwareRouter.route_request lookup -> choice -> insert
//       (src/vllm_router/routers/routing_logic.py:414-423)
//
// Kernel pipeline of one routed batch (all on one stream, no host sync):
//   k_prep_sums / k_prep_scan   chunk counts -> CSR chunk_base, tile map
//   [k_count_chars, k_utf8_bounds   only for batches that are not pure ASCII]
//   k_hash_chunks               XXH64 of every chunk, one thread per chunk,
//                               bytes staged by cp.async into swizzled smem
//   k_claim                     path keys (warp scan) + find-or-create in the
//                               open-addressing table + "first toucher" atomicMax
//   k_select                    per request: deepest usable node, root draw or leader
//   k_commit                    resolve leader chains, OR the endpoint into new nodes
#include <cstdarg>
#include <mutex>
#include <algorithm>
#include <vector>

#include "common.cuh"
#define B200R_TH_FN __device__ __forceinline__
#define B200R_TH_MUL(C, x) ::b200r::mulc<C>(x)
#define B200R_TH_ROTL(R, x) ::b200r::rotl<R>(x)
#include "tailhash.cuh"

namespace b200r {

// ------------------------------------------------------------------- errors
static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char *what)
{
    set_error("%s: %s", what, cudaGetErrorString(e));
    return e == cudaErrorMemoryAllocation ? B200R_E_NOMEM : B200R_E_CUDA;
}

// ------------------------------------------------------------- profiling
// Optional per-kernel timing with CUDA events on the launching stream
// (b200r_prof_enable / b200r_prof_read).  Off by default: the hot path then
// records nothing.  g_launches counts every kernel this library enqueues.
enum { PK_PREP = 0, PK_UTF8, PK_HASH, PK_CLAIM, PK_SELECT, PK_COMMIT, PK_LOOKUP, PK_INSERT,
       PK_CLASSIFY, PK_PUSH, PK_WAIT, PK_APPLY, PK_RESOLVE, PK_COUNT };
struct ProfSpan { int kid; cudaEvent_t a, b; };
static bool g_prof_on = false;
static std::vector<ProfSpan> g_spans;
static long long g_launches = 0;
static bool g_claim_by_tiles = true;    // route stage: claim per 32-chunk tile (else per request)

struct ProfScope {
    bool on; cudaStream_t st; ProfSpan sp;
    ProfScope(int kid, cudaStream_t s) : on(g_prof_on), st(s)
    {
        if (!on) return;
        sp.kid = kid;
        cudaEventCreate(&sp.a); cudaEventCreate(&sp.b);
        cudaEventRecord(sp.a, st);
    }
    ~ProfScope()
    {
        if (!on) return;
        cudaEventRecord(sp.b, st);
        g_spans.push_back(sp);
    }
};

constexpr int PREP_BLOCK = 256;    // requests per CTA of the prep kernels
constexpr int WARPS_PER_CTA = 8;   // per-request kernels: one warp per request

// =========================================================================
// Prep: per-request chunk counts -> exclusive scan (CSR) + tile -> request map
// =========================================================================
__device__ __forceinline__ int chunks_of(const int32_t *lens, const int32_t *nchars, int i, int cc)
{
    int nc = nchars ? nchars[i] : lens[i];
    return (nc + cc - 1) / cc;
}

__device__ __forceinline__ int block_sum_1024(int v, int *s_red)
{   // all PREP_BLOCK threads get the block total
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_red[w] = v;
    __syncthreads();
    int t = lane < PREP_BLOCK / 32 ? s_red[lane] : 0;
    for (int o = 16; o; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    __syncthreads();
    return t;
}

__global__ void __launch_bounds__(PREP_BLOCK)
k_prep_sums(const int32_t *__restrict__ lens, const int32_t *__restrict__ nchars, int n, int cc,
            int32_t *__restrict__ block_sums, const int32_t *__restrict__ n_dev,
            unsigned int *__restrict__ ticket)
{
    __shared__ int s_red[32];
    if (n_dev) n = *n_dev;          // batch size known only on the device (partitioned multi-GPU)
    if (ticket && blockIdx.x == 0 && threadIdx.x == 0) *ticket = 0;   // for the fused kernel
    int i = blockIdx.x * PREP_BLOCK + threadIdx.x;
    int c = (i < n) ? chunks_of(lens, nchars, i, cc) : 0;
    int tot = block_sum_1024(c, s_red);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(PREP_BLOCK)
k_prep_scan(const int32_t *__restrict__ lens, const int32_t *__restrict__ nchars, int n, int cc,
            const int32_t *__restrict__ block_sums, int32_t *__restrict__ chunk_base,
            int32_t *__restrict__ tile_first, int64_t max_chunks, TrieDev *dev,
            const uint8_t *__restrict__ bytes, const int64_t *__restrict__ starts,
            uint64_t *__restrict__ hashes, const int32_t *__restrict__ n_dev,
            unsigned int *__restrict__ lb_flag)
{
    __shared__ int s_red[32];
    __shared__ int s_wsum[32];
    if (n_dev) {
        n = *n_dev;
        if (n == 0 && blockIdx.x == 0 && threadIdx.x == 0) chunk_base[0] = 0;
    }
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    // offset of this block = sum of the sums of all earlier blocks
    int part = 0;
    for (int b = tid; b < (int)blockIdx.x; b += PREP_BLOCK) part += block_sums[b];
    const int offset = block_sum_1024(part, s_red);

    const int i = blockIdx.x * PREP_BLOCK + tid;
    const int c = (i < n) ? chunks_of(lens, nchars, i, cc) : 0;
    int inc = c;                                    // inclusive warp scan
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_wsum[w] = inc;
    __syncthreads();
    int wbase = 0;
    for (int k = 0; k < w; k++) wbase += s_wsum[k];
    const int base = offset + wbase + inc - c;      // exclusive prefix
    if (i < n) {
        chunk_base[i] = base;
        if (i == n - 1) {
            chunk_base[n] = base + c;
            if (dev && (int64_t)(base + c) > max_chunks) atomicOr(&dev->status, B200R_S_CHUNK_CAP);
        }
        // warp tiles (32 chunks) whose first chunk belongs to this request
        if (c > 0) {
            int t0 = (base + 31) / 32;
            int64_t lim = (int64_t)base + c;
            if (lim > max_chunks) lim = max_chunks;
            for (int t = t0; (int64_t)t * 32 < lim; t++) {
                tile_first[t] = i;
                if (lb_flag) lb_flag[t] = 0;          // look-back state of the fused kernel
            }
            // The ragged last chunk of an ASCII request is hashed here, one thread per
            // request, so that k_hash_chunks is left with full 128-byte chunks only.
            const int ln = lens[i];
            const int nc = nchars ? nchars[i] : ln;
            const int tail = ln - (c - 1) * cc;
            const int64_t gl = (int64_t)base + c - 1;
            if (hashes && nc == ln && tail < cc && gl < max_chunks) {
                const uint8_t *q = bytes + starts[i];
                if ((((uintptr_t)q) & 15) == 0) {
                    const uint8_t *tp = q + (ln - tail);
                    if (tail < 128 && (((uintptr_t)tp) & 15) == 0) {
                        // fetch the whole tail with independent 16-byte loads first (one DRAM
                        // round trip instead of one per stripe), then hash it out of registers
                        // (tailhash.cuh: static register indexing, no local memory)
                        uint64_t wv[16];
#pragma unroll
                        for (int u = 0; u < 8; u++) {
                            const uint4 q = (u * 16 < tail) ? __ldg((const uint4 *)tp + u) : make_uint4(0, 0, 0, 0);
                            wv[2 * u] = u64_of(q.x, q.y);
                            wv[2 * u + 1] = u64_of(q.z, q.w);
                        }
                        hashes[gl] = b200r_th::xxh64_tail_words(wv, tail);
                    } else {
                        hashes[gl] = xxh64_global(tp, tail);
                    }
                }
            }
        }
    }
}

// code points of each request (only when the caller did not supply nchars)
__global__ void k_count_chars(const uint8_t *__restrict__ bytes, const int64_t *__restrict__ starts,
                              const int32_t *__restrict__ lens, int n, int32_t *__restrict__ nchars)
{
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= n) return;
    const uint8_t *p = bytes + starts[warp];
    const int len = lens[warp];
    int cnt = 0;
    for (int o = lane; o < len; o += 32) cnt += ((p[o] & 0xC0) != 0x80);
    for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if (lane == 0) nchars[warp] = cnt;
}

// byte offset (inside the request) of every chunk start, for non-ASCII requests
__global__ void k_utf8_bounds(const uint8_t *__restrict__ bytes, const int64_t *__restrict__ starts,
                              const int32_t *__restrict__ lens, const int32_t *__restrict__ nchars,
                              const int32_t *__restrict__ chunk_base, int n, int cc,
                              int64_t max_chunks, int32_t *__restrict__ chunk_off)
{
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= n) return;
    const int len = lens[warp];
    if (nchars[warp] == len) return;             // ASCII: offsets are arithmetic
    const uint8_t *p = bytes + starts[warp];
    const int b0 = chunk_base[warp];
    int run = 0;                                  // code points before this block
    for (int o = 0; o < len; o += 32) {
        int pos = o + lane;
        bool lead = pos < len && ((p[pos] & 0xC0) != 0x80);
        unsigned bal = __ballot_sync(0xffffffffu, lead);
        int idx = run + __popc(bal & ((1u << lane) - 1));
        if (lead && idx % cc == 0) {
            int64_t g = (int64_t)b0 + idx / cc;
            if (g < max_chunks) chunk_off[g] = pos;
        }
        run += __popc(bal);
    }
}

// =========================================================================
// Chunk hashing
// =========================================================================
__device__ __forceinline__ void cp_async16(uint32_t smem_dst, const void *src, int src_bytes)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(smem_dst), "l"(src),
                 "r"(src_bytes)
                 : "memory");
}
// the same with an L2 evict-first hint: prompt bytes are read exactly once, so they should not
// push the table entries and scratch arrays of the route stage (running on the other stream)
// out of L2
__device__ __forceinline__ uint64_t l2_evict_first_policy()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void cp_async16_stream(uint32_t smem_dst, const void *src, uint64_t pol)
{
    asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;\n" ::"r"(smem_dst), "l"(src),
                 "l"(pol)
                 : "memory");
}
__device__ __forceinline__ void cp_async_wait_all()
{
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
}

__device__ __forceinline__ uint4 lds128(uint32_t saddr)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr));
    return v;
}

// Staged chunk rows: row c of the tile is 128 bytes at tile + 128*c whose eight
// 16-byte units are XOR-swizzled by (c & 7), so the LDS.128 reads of the 8
// threads of one shared-memory phase (stride 128 B) hit 8 different bank groups.
//
// XXH64 of a full 128-byte chunk: straight-line, all eight LDS.128 up front.
__device__ __forceinline__ uint64_t xxh64_row128(uint32_t row, int sw)
{
    uint4 q[8];
#pragma unroll
    for (int u = 0; u < 8; u++) q[u] = lds128(row + ((u ^ sw) << 4));
    uint64_t v1 = XP1 + XP2, v2 = XP2, v3 = 0, v4 = 0 - XP1;
#pragma unroll
    for (int s = 0; s < 4; s++) {
        v1 = xround(v1, join64(q[2 * s].x, q[2 * s].y));
        v2 = xround(v2, join64(q[2 * s].z, q[2 * s].w));
        v3 = xround(v3, join64(q[2 * s + 1].x, q[2 * s + 1].y));
        v4 = xround(v4, join64(q[2 * s + 1].z, q[2 * s + 1].w));
    }
    uint64_t h = rotl<1>(v1) + rotl<7>(v2) + rotl<12>(v3) + rotl<18>(v4);
    h = xmerge(h, v1); h = xmerge(h, v2); h = xmerge(h, v3); h = xmerge(h, v4);
    h += 128;
    return xavalanche(h);
}

// XXH64 of a staged chunk of any length 1..128 (ragged last chunk of a request).
__device__ __noinline__ uint64_t xxh64_row_any(const uint8_t *row, int sw, int len)
{
#define ROW_OFF(o) (((((o) >> 4) ^ sw) << 4) | ((o)&15))
    uint64_t h;
    const int nst = len >> 5;
    if (nst) {
        uint64_t v1 = XP1 + XP2, v2 = XP2, v3 = 0, v4 = 0 - XP1;
        for (int s = 0; s < nst; s++) {
            const uint4 a = *(const uint4 *)(row + (((2 * s) ^ sw) << 4));
            const uint4 b = *(const uint4 *)(row + (((2 * s + 1) ^ sw) << 4));
            v1 = xround(v1, join64(a.x, a.y));
            v2 = xround(v2, join64(a.z, a.w));
            v3 = xround(v3, join64(b.x, b.y));
            v4 = xround(v4, join64(b.z, b.w));
        }
        h = rotl<1>(v1) + rotl<7>(v2) + rotl<12>(v3) + rotl<18>(v4);
        h = xmerge(h, v1); h = xmerge(h, v2); h = xmerge(h, v3); h = xmerge(h, v4);
    } else {
        h = XP5;
    }
    h += (uint64_t)len;
    int o = nst << 5;
    // 8- and 4-byte steps never straddle a 16-byte unit (o is a multiple of 8 / 4)
    while (o + 8 <= len) {
        const uint64_t w = *(const uint64_t *)(row + ROW_OFF(o));
        h ^= xround(0, w);
        h = mulc<XP1>(rotl<27>(h)) + XP4;
        o += 8;
    }
    if (o + 4 <= len) {
        const uint32_t w = *(const uint32_t *)(row + ROW_OFF(o));
        h ^= mulc<XP1>((uint64_t)w);
        h = mulc<XP2>(rotl<23>(h)) + XP3;
        o += 4;
    }
    while (o < len) {
        h ^= mulc<XP5>((uint64_t)row[ROW_OFF(o)]);
        h = mulc<XP1>(rotl<11>(h));
        o++;
    }
    return xavalanche(h);
#undef ROW_OFF
}

// ---------------------------------------------------------------------------
// k_hash_chunks: one thread per chunk, and every WARP is an independent worker
// on its own tile of 32 consecutive chunks -- no block-wide barrier anywhere, so
// the 48 resident warps of an SM (6 CTAs x 8 warps, 4 KB of staging each) are
// 48 independent latency chains and ~190 KB of prompt bytes stay in flight.
//   map     tile -> requests with two rounds of independent global loads:
//           tile_first[tile], then lane j fetches everything about request
//           r0 + j; head positions are combined with a warp OR-reduction and
//           each chunk lane pulls its request's record by shuffle.
//   stage   the warp copies its 32 chunks with cp.async (coalesced: one
//           instruction moves 4 chunks x 128 B) into XOR-swizzled shared memory
//   hash    full 128-byte chunks: straight-line XXH64, eight LDS.128 up front.
// Ragged last chunks of ASCII requests are not touched here: k_prep_scan (one
// thread per request) hashes them, so this kernel has a single code path.
// Unaligned or multi-byte-UTF-8 chunks take the generic global-memory routine.
constexpr int WTILE = 32;                 // chunks per warp tile
constexpr int HASH_WARPS = 8;             // warps per CTA

struct ReqRec { int b, b1, ln, nc; long long st; };

// Arguments of the fused variant (hash + path keys + claim in one kernel).
struct FuseArgs {
    Entry *tab; uint64_t *hs; int cap_log2; TrieDev *dev;
    uint64_t avail; unsigned long long tag;
    uint32_t *slots;
    int32_t *chunk_req;              // request owning each chunk: written by MODE 0, read by MODE 2
    unsigned int *ticket;            // CTA ticket counter (zeroed by k_prep_sums)
    unsigned int *lb_flag;           // per warp tile: 0 nothing, 1 partial key published, 2 final
    uint64_t *lb_part, *lb_final;    // per warp tile: {k1, stamp, k2, stamp}
};
constexpr uint32_t PREOK_BIT = 0x80000000u;

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u32(unsigned *p, unsigned v)
{
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_cg_2u64(uint64_t *p, uint64_t a, uint64_t b)
{
    asm volatile("st.global.cg.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory");
}
__device__ __forceinline__ void ld_cg_2u64(const uint64_t *p, uint64_t &a, uint64_t &b)
{
    asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}

// A^n for n = 1..32, both key lanes (A^(2^k) are compile-time constants)
__device__ __forceinline__ void key_pow(int n, uint64_t &r1, uint64_t &r2)
{
    r1 = 1; r2 = 1;
#define B200R_POW_STEP(BIT)                                                     \
    if (n & BIT) { r1 = mulc<cpow(KA1, BIT)>(r1); r2 = mulc<cpow(KA2, BIT)>(r2); }
    B200R_POW_STEP(1) B200R_POW_STEP(2) B200R_POW_STEP(4) B200R_POW_STEP(8) B200R_POW_STEP(16)
    B200R_POW_STEP(32)
#undef B200R_POW_STEP
}

// CLAIM = false: chunk hashing only (a1, and the hash stage of the staged paths).
// CLAIM = true : fused single-GPU path -- after hashing, the warp computes the path keys
//   of its 32 chunks (segmented Kogge-Stone scan; a request that started in an earlier
//   tile gets its carry by decoupled look-back over per-tile published keys, CTAs take
//   tickets so that a waited-for tile is always running or done), then finds-or-creates
//   every node and classifies it exactly like k_claim.  The chunk hashes never go to
//   memory and the table latency of one warp overlaps the hashing of the others.
//   MODE 0  chunk hashing only (a1, and the hash stage of the staged paths)
//   MODE 1  fused: hash + path keys + claim (carry between tiles by decoupled look-back)
//   MODE 2  claim from hashes already in memory, one warp per 32-chunk tile: every lane
//           works (a warp per request leaves a third of them idle on ~21-chunk prompts), no
//           staging buffer, and the carry of a request that started in an earlier tile is
//           recomputed from that request's earlier hashes (no communication between tiles)
template <int MODE>
__global__ void __launch_bounds__(HASH_WARPS * 32, MODE == 2 ? 8 : 6)
k_hash_tiles(const uint8_t *__restrict__ bytes, const int64_t *__restrict__ starts,
             const int32_t *__restrict__ lens, const int32_t *__restrict__ nchars,
             const int32_t *__restrict__ chunk_base, const int32_t *__restrict__ tile_first,
             const int32_t *__restrict__ chunk_off, int n, int cc, int64_t max_chunks,
             uint64_t *__restrict__ hashes, FuseArgs fa)
{
    constexpr bool CLAIM = MODE != 0;
    __shared__ __align__(1024) uint8_t s_tile[MODE == 2 ? 16 : HASH_WARPS * WTILE * 128];
    __shared__ unsigned s_bid;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    unsigned bid = blockIdx.x;
    if (MODE == 1) {
        if (threadIdx.x == 0) s_bid = atomicAdd(fa.ticket, 1u);
        __syncthreads();
        bid = s_bid;
    }
    const int64_t tile = (int64_t)bid * HASH_WARPS + w;
    const int64_t G0 = tile * WTILE;
    int64_t total = chunk_base[n];
    if (total > max_chunks) total = max_chunks;
    if (G0 >= total) return;                                   // whole warp

    // ---- map: which request owns each of the 32 chunks?  (MODE 2 reads what MODE 0 found)
    ReqRec mine = { 0, 0, 0, 0, 0 };      // record of the request that owns this lane's chunk
    int mine_r = 0;
    if (MODE == 2) {
        if (G0 + lane < total) {
            mine_r = fa.chunk_req[G0 + lane];
            mine.b = chunk_base[mine_r];
        }
    }
    const int r0 = MODE == 2 ? 0 : tile_first[tile];
    if (MODE != 2) mine_r = r0;
    for (int jb = 0; MODE != 2; jb += 32) {
        const int r = r0 + jb + lane;
        ReqRec q = { 0, 0, 0, 0, 0 };
        bool live = false;
        if (r < n) {
            q.b = chunk_base[r]; q.b1 = chunk_base[r + 1];
            if (MODE != 2) {                 // MODE 2 works from hashes: no prompt bytes needed
                q.st = starts[r]; q.ln = lens[r];
                q.nc = nchars ? nchars[r] : q.ln;
            }
            live = (int64_t)q.b < G0 + WTILE;
        }
        const bool nonempty = live && q.b1 > q.b;
        const int pos = (int64_t)q.b > G0 ? (int)(q.b - G0) : 0;
        const unsigned heads = __reduce_or_sync(0xffffffffu, nonempty ? (1u << pos) : 0u);
        const unsigned ne = __ballot_sync(0xffffffffu, nonempty);
        const unsigned below = heads & (0xffffffffu >> (31 - lane));   // heads at positions <= lane
        // lanes with `below` != 0: a request of this block starts at or before my chunk
        int src = 0;
        if (below) {
            const int hp = 31 - __clz(below);
            const int k = __popc(heads & ((1u << hp) - 1));            // its rank among non-empty ones
            src = (ne == (0xffffffffu >> (32 - __popc(ne)))) ? k : (int)__fns(ne, 0, k + 1);
        }
        const int tb = __shfl_sync(0xffffffffu, q.b, src), tb1 = __shfl_sync(0xffffffffu, q.b1, src);
        const int tln = __shfl_sync(0xffffffffu, q.ln, src), tnc = __shfl_sync(0xffffffffu, q.nc, src);
        const long long tst = __shfl_sync(0xffffffffu, q.st, src);
        if (below) {
            mine.b = tb; mine.b1 = tb1; mine.ln = tln; mine.nc = tnc; mine.st = tst;
            mine_r = r0 + jb + src;
        }
        // more requests can start inside the tile only if the last lane's request ends inside it
        const int last_b1 = __shfl_sync(0xffffffffu, q.b1, 31);
        const bool last_live = __shfl_sync(0xffffffffu, (int)live, 31);
        if (!(last_live && (int64_t)last_b1 < G0 + WTILE && r0 + jb + 32 < n)) break;
    }

    // ---- chunk descriptor
    const int64_t g = G0 + lane;
    const bool valid = g < total;
    // chunk -> request map for the tile claim kernel (stored now: mine_r is not needed again)
    if (MODE == 0 && valid && fa.chunk_req) fa.chunk_req[g] = mine_r;
    const uint8_t *p = nullptr;
    int clen = 0, kind = 0;               // 0 none / done elsewhere, 1 full staged, 3 generic
    int d = 0;                            // depth of the chunk inside its request (0-based)
    if (valid) d = (int)(g - mine.b);
    if (valid && MODE != 2) {
        int off;
        if (mine.nc == mine.ln) {                        // ASCII: chunk = cc bytes
            off = d * cc;
            clen = min(cc, mine.ln - off);
        } else {                                         // code-point boundaries (k_utf8_bounds)
            off = chunk_off[g];
            const int end = (g + 1 < (int64_t)mine.b1 && g + 1 < max_chunks) ? chunk_off[g + 1] : mine.ln;
            clen = end - off;
        }
        p = bytes + mine.st + off;
        const bool aligned = (((uintptr_t)p & 15) == 0);
        const bool tail_done_by_prep = (mine.nc == mine.ln) && g + 1 == (int64_t)mine.b1 &&
                                       clen < cc && ((((uintptr_t)bytes + mine.st) & 15) == 0);
        kind = tail_done_by_prep ? 0 : ((aligned && clen == 128) ? 1 : 3);
    }

    // ---- stage this warp's full chunks, then hash (MODE 2: the hashes are in memory already)
    uint64_t h = 0;
    if (MODE == 2) {
        if (valid) h = hashes[g];
    } else {
        const uint32_t wrow = (uint32_t)__cvta_generic_to_shared(s_tile) + (uint32_t)(w * WTILE) * 128;
        const unsigned long long my_src = kind == 1 ? (unsigned long long)(uintptr_t)p : 0ULL;
        // lane copies unit u of chunk c = 4k + sub; (c & 7) = sub ^ 4(k&1), so the swizzled
        // destination is one of two per-lane bases plus k * 512
        const int u = lane & 7, sub = lane >> 3;
        const uint64_t pol = l2_evict_first_policy();
        uint32_t d_even = wrow + sub * 128 + ((u ^ sub) << 4);
        asm volatile("" : "+r"(d_even));
        const uint32_t d_odd = d_even ^ 64u;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const unsigned long long src = __shfl_sync(0xffffffffu, my_src, 4 * k + sub);
            if (src)
                cp_async16_stream(((k & 1) ? d_odd : d_even) + k * 512,
                                  (const uint8_t *)(uintptr_t)src + (u << 4), pol);
        }
        // the ragged last chunk of a request was hashed by k_prep_scan: fetch it meanwhile
        if (MODE == 1 && valid && kind == 0) h = hashes[g];
        cp_async_wait_all();
        __syncwarp();
        if (kind == 1) h = xxh64_row128(wrow + lane * 128, lane & 7);
        else if (kind == 3) h = xxh64_global(p, clen);
    }
    if (!CLAIM) {
        if (kind) hashes[g] = h;
        return;
    }

    // ---- path keys: segmented scan over the tile
    const bool inhead = valid && d > lane;          // request started in an earlier tile
    const int dseg = !valid ? 0 : (inhead ? lane : d);
    uint64_t p1 = h, p2 = rotl<32>(h);
    if (valid && d == 0) { p1 += KA1 * KROOT1; p2 += KA2 * KROOT2; }
#define B200R_SEG_STEP(S)                                          \
    {                                                              \
        const uint64_t t1 = __shfl_up_sync(0xffffffffu, p1, S);    \
        const uint64_t t2 = __shfl_up_sync(0xffffffffu, p2, S);    \
        constexpr uint64_t m1 = cpow(KA1, S), m2 = cpow(KA2, S);   \
        if (dseg >= S) { p1 += m1 * t1; p2 += m2 * t2; }           \
    }
    B200R_SEG_STEP(1) B200R_SEG_STEP(2) B200R_SEG_STEP(4) B200R_SEG_STEP(8) B200R_SEG_STEP(16)
#undef B200R_SEG_STEP

    // Publish the key at the end of this tile for the next tile's carry.  Each key half
    // travels in its own 16-byte word {half, stamp} written with one vector store, and a
    // reader accepts a word only if its stamp is the one of this batch: no fence, no
    // acquire (those would invalidate L1, which serves the hot table entries), and no
    // clearing between batches.
    const bool head_tile = __shfl_sync(0xffffffffu, (int)inhead, 0);    // lane 0 continues a request
    const bool last_valid = __shfl_sync(0xffffffffu, (int)valid, 31);
    const bool last_inhead = __shfl_sync(0xffffffffu, (int)inhead, 31); // whole tile inside one request
    const uint64_t stamp_part = fa.tag | 1u, stamp_final = fa.tag | 2u;
    if (MODE == 1 && lane == 31 && last_valid) {
        if (last_inhead) { st_cg_2u64(fa.lb_part + 4 * tile, p1, stamp_part); st_cg_2u64(fa.lb_part + 4 * tile + 2, p2, stamp_part); }
        else { st_cg_2u64(fa.lb_final + 4 * tile, p1, stamp_final); st_cg_2u64(fa.lb_final + 4 * tile + 2, p2, stamp_final); }
    }
    if (MODE == 2 && head_tile) {
        // the request of lane 0 started d0 chunks before this tile: recompute the key at its
        // depth d0 from its earlier hashes (they are in memory; typically one warp load)
        const int d0 = __shfl_sync(0xffffffffu, d, 0);
        const int b_req = __shfl_sync(0xffffffffu, mine.b, 0);
        Key carry = { KROOT1, KROOT2 };
        for (int pb = 0; pb < d0; pb += 32) {
            const int x = pb + lane;
            const uint64_t hh = x < d0 ? hashes[b_req + x] : 0;
            const Key k = key_scan_warp(hh, carry, lane);
            const int last = min(31, d0 - 1 - pb);
            carry.k1 = __shfl_sync(0xffffffffu, k.k1, last);
            carry.k2 = __shfl_sync(0xffffffffu, k.k2, last);
        }
        if (inhead) {
            const ulonglong2 w = g_key_pow[lane];
            p1 += w.x * carry.k1; p2 += w.y * carry.k2;
        }
    }
    if (MODE == 1 && head_tile) {
        // decoupled look-back: K_in = P(T-1) + A^32 P(T-2) + ... + A^(32m) F(j)
        uint64_t a1 = 0, a2 = 0, m1 = 1, m2 = 1;
        for (int64_t j = tile - 1; j >= 0; j--) {
            uint64_t x1 = 0, x2 = 0, s1, s2;
            int got = 0;
            for (long spins = 0; spins < 40000000L && !got; spins++) {
                ld_cg_2u64(fa.lb_final + 4 * j, x1, s1);
                ld_cg_2u64(fa.lb_final + 4 * j + 2, x2, s2);
                if (s1 == stamp_final && s2 == stamp_final) { got = 2; break; }
                ld_cg_2u64(fa.lb_part + 4 * j, x1, s1);
                ld_cg_2u64(fa.lb_part + 4 * j + 2, x2, s2);
                if (s1 == stamp_part && s2 == stamp_part) { got = 1; break; }
                __nanosleep(32);
            }
            if (!got) { if (lane == 0) atomicOr(&fa.dev->status, B200R_S_LOOKBACK); break; }
            a1 += m1 * x1; a2 += m2 * x2;
            if (got == 2) break;
            m1 = mulc<cpow(KA1, 32)>(m1); m2 = mulc<cpow(KA2, 32)>(m2);
        }
        if (inhead) {
            uint64_t w1, w2;
            key_pow(lane + 1, w1, w2);
            p1 += w1 * a1; p2 += w2 * a2;
        }
        if (lane == 31 && last_valid && last_inhead) {
            st_cg_2u64(fa.lb_final + 4 * tile, p1, stamp_final);
            st_cg_2u64(fa.lb_final + 4 * tile + 2, p2, stamp_final);
        }
    }

    // ---- find-or-create + classify (same as k_claim)
    if (valid) {
        bool created = false;
        uint64_t m = 0, cur = 0;
        uint32_t s;
        if (MODE == 2) {
            s = table_claim_mf(fa.tab, fa.cap_log2, p1, p2, h, &created, &m, &cur);
        } else {
            s = table_claim(fa.tab, fa.hs, fa.cap_log2, p1, p2, h, &created);
            if (s != SLOT_NONE && !created) m = fa.tab[s].mask;
        }
        if (s == SLOT_NONE) {
            atomicOr(&fa.dev->status, B200R_S_TABLE_FULL);
            fa.slots[g] = SLOT_NONE;
        } else if (m & fa.avail) {
            fa.slots[g] = s | PREOK_BIT;
        } else {
            fa.slots[g] = s;
            const unsigned long long minev = fa.tag | (unsigned long long)(~(unsigned)mine_r);
            if (MODE != 2 && !created) cur = *(volatile unsigned long long *)&fa.tab[s].first;
            if (cur < minev) atomicMax(&fa.tab[s].first, minev);
        }
    }
}

// =========================================================================
// Table kernels (one warp per request, lanes = chunk depths)
// =========================================================================
struct ReqSpan { int b0, C; };

__device__ __forceinline__ ReqSpan req_span(const int32_t *chunk_base, int i, int64_t max_chunks)
{
    ReqSpan s;
    s.b0 = chunk_base[i];
    int64_t e = chunk_base[i + 1];
    if (e > max_chunks) e = max_chunks;
    s.C = (e > s.b0) ? (int)(e - s.b0) : 0;
    return s;
}

// a2: frozen-table longest prefix match.  All depths of a 32-chunk block are
// probed in parallel; the first miss is located with __ballot_sync/__ffs.
// Relies on the trie invariant child.endpoints subset-of parent.endpoints
// (inserts add the endpoint to every node of a path), so the reference's
// running intersection equals mask(node) & avail.
__global__ void __launch_bounds__(WARPS_PER_CTA * 32)
k_lookup(const Entry *__restrict__ tab, int cap_log2, const uint64_t *__restrict__ hashes,
         const int32_t *__restrict__ chunk_base, int n, uint64_t avail,
         int32_t *__restrict__ depth_out, uint64_t *__restrict__ matched_out)
{
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (i >= n) return;
    const ReqSpan sp = req_span(chunk_base, i, INT64_MAX);
    Key carry = { KROOT1, KROOT2 };
    int depth = 0;
    uint64_t sel = avail;
    for (int blk = 0; blk * 32 < sp.C; blk++) {
        const int d = blk * 32 + lane;
        const bool valid = d < sp.C;
        const uint64_t h = valid ? hashes[sp.b0 + d] : 0;
        const Key k = key_scan_warp(h, carry, lane);
        carry.k1 = __shfl_sync(0xffffffffu, k.k1, 31);
        carry.k2 = __shfl_sync(0xffffffffu, k.k2, 31);
        uint64_t m = 0;
        if (valid) {
            uint32_t s = table_find(tab, cap_log2, k.k1, k.k2);
            if (s != SLOT_NONE) m = tab[s].mask & avail;
        }
        const unsigned bal = __ballot_sync(0xffffffffu, m != 0);
        const int nok = (bal == 0xffffffffu) ? 32 : (__ffs(~bal) - 1);
        if (nok > 0) {
            sel = __shfl_sync(0xffffffffu, m, nok - 1);
            depth += nok;
        }
        if (nok < 32) break;
    }
    if (lane == 0) {
        depth_out[i] = depth;
        matched_out[i] = sel;
    }
}

// a3: order-independent batched insert.
__global__ void __launch_bounds__(WARPS_PER_CTA * 32)
k_insert(Entry *__restrict__ tab, uint64_t *__restrict__ hs, int cap_log2, TrieDev *dev,
         const uint64_t *__restrict__ hashes,
         const int32_t *__restrict__ chunk_base, int n, const int32_t *__restrict__ endpoint)
{
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (i >= n) return;
    const int e = endpoint[i];
    if (e < 0 || e > 63) return;
    const uint64_t bit = 1ULL << e;
    if (lane == 0 && !(dev->root_mask & bit)) atomicOr((unsigned long long *)&dev->root_mask, bit);
    const ReqSpan sp = req_span(chunk_base, i, INT64_MAX);
    Key carry = { KROOT1, KROOT2 };
    for (int blk = 0; blk * 32 < sp.C; blk++) {
        const int d = blk * 32 + lane;
        const bool valid = d < sp.C;
        const uint64_t h = valid ? hashes[sp.b0 + d] : 0;
        const Key k = key_scan_warp(h, carry, lane);
        carry.k1 = __shfl_sync(0xffffffffu, k.k1, 31);
        carry.k2 = __shfl_sync(0xffffffffu, k.k2, 31);
        bool created = false;
        if (valid) {
            uint32_t s = table_claim(tab, hs, cap_log2, k.k1, k.k2, h, &created);
            if (s == SLOT_NONE) atomicOr(&dev->status, B200R_S_TABLE_FULL);
            else if (created || !(tab[s].mask & bit))
                atomicOr((unsigned long long *)&tab[s].mask, bit);
        }
    }
}

// a5 step 1: claim every node of every path; classify against the pre-batch
// masks; record who touches a not-yet-usable node first.
//   slots[g] = slot | PREOK_BIT   node usable before the batch (mask & avail != 0)
//   slots[g] = slot               otherwise; entry.first = (epoch, ~min request index)


// Per-rank device-resident bookkeeping of the peer-exchange (P2P) multi-GPU path.
constexpr int P2P_MAX_WORLD = 64;
struct P2PDyn {
    unsigned long long seq_base;     // global sequence number of the next batch's request 0
    uint32_t my_off, total_n;        // position of this shard / size of the current global batch
    int32_t meta[2 * P2P_MAX_WORLD]; // {record count, shard size} of every rank, current batch
};

__device__ __forceinline__ void
claim_request(int i, int lane, Entry *__restrict__ tab, uint64_t *__restrict__ hs, int cap_log2,
              TrieDev *dev, const uint64_t *__restrict__ hashes,
              const int32_t *__restrict__ chunk_base, int64_t max_chunks, uint64_t avail,
              unsigned long long tag, uint32_t *__restrict__ slots)
{
    const ReqSpan sp = req_span(chunk_base, i, max_chunks);
    Key carry = { KROOT1, KROOT2 };
    for (int blk = 0; blk * 32 < sp.C; blk++) {
        const int d = blk * 32 + lane;
        const bool valid = d < sp.C;
        const uint64_t h = valid ? hashes[sp.b0 + d] : 0;
        const Key k = key_scan_warp(h, carry, lane);
        carry.k1 = __shfl_sync(0xffffffffu, k.k1, 31);
        carry.k2 = __shfl_sync(0xffffffffu, k.k2, 31);
        bool created = false;
        if (valid) {
            uint32_t s = table_claim(tab, hs, cap_log2, k.k1, k.k2, h, &created);
            if (s == SLOT_NONE) {
                atomicOr(&dev->status, B200R_S_TABLE_FULL);
                slots[sp.b0 + d] = SLOT_NONE;
            } else {
                // masks do not change during this kernel: L1 may serve them (hot nodes)
                const uint64_t m = created ? 0 : tab[s].mask;
                if (m & avail) {
                    slots[sp.b0 + d] = s | PREOK_BIT;
                } else {
                    slots[sp.b0 + d] = s;
                    // hot nodes are touched by many requests of a batch: skip the atomic when
                    // an earlier request of this batch is already recorded (a stale read only
                    // costs a redundant atomicMax)
                    const unsigned long long mine = tag | (unsigned long long)(~(unsigned)i);
                    unsigned long long cur = 0;
                    if (!created) cur = *(volatile unsigned long long *)&tab[s].first;
                    if (cur < mine) atomicMax(&tab[s].first, mine);
                }
            }
        }
    }
}

__global__ void __launch_bounds__(WARPS_PER_CTA * 32)
k_claim(Entry *__restrict__ tab, uint64_t *__restrict__ hs, int cap_log2, TrieDev *dev,
        const uint64_t *__restrict__ hashes, const int32_t *__restrict__ chunk_base, int n,
        int64_t max_chunks, uint64_t avail, unsigned long long tag, uint32_t *__restrict__ slots,
        const int32_t *__restrict__ n_dev)
{
    // grid-stride over requests: with n_dev the batch size is only known on the device and
    // the grid is an estimate
    const int lane = threadIdx.x & 31;
    if (n_dev) n = *n_dev;
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += (gridDim.x * blockDim.x) >> 5)
        claim_request(i, lane, tab, hs, cap_log2, dev, hashes, chunk_base, max_chunks, avail, tag, slots);
}

// a5 step 2: deepest node usable by request i = last depth d with
//   PREOK(node) or first(node) < i      (monotone in d)
// If that node was usable before the batch (or depth 0) the request is a root
// and draws from its candidate set; else it follows the request that touched
// the node first.
__device__ __forceinline__ void
select_request(int i, const Entry *__restrict__ tab, const int32_t *__restrict__ chunk_base,
               int64_t max_chunks, const uint32_t *__restrict__ slots, uint64_t avail,
               unsigned long long tag, uint64_t seed, uint64_t seq0, int32_t *__restrict__ leader,
               int32_t *__restrict__ choice, int32_t *__restrict__ depth_out,
               uint64_t *__restrict__ matched_out, uint32_t req_off, const uint32_t *__restrict__ pos)
{
    const ReqSpan sp = req_span(chunk_base, i, max_chunks);
    const uint32_t *sl = slots + sp.b0;
    int dstar = 0;
    uint32_t star_slot = SLOT_NONE, star_first = 0;
    bool go = true;
    // slots are fetched eight at a time (independent loads: one L2 round trip per group); the
    // first-toucher words stay lazy -- the walk usually ends at the first of them, and each
    // is a random table sector
    for (int d0 = 0; d0 < sp.C && go; d0 += 8) {
        uint32_t v[8];
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = (d0 + k < sp.C) ? sl[d0 + k] : SLOT_NONE;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            if (!go) break;
            const uint32_t x = v[k];
            bool ok = false;
            uint32_t f = 0;
            if (x != SLOT_NONE) {
                if (x & PREOK_BIT) ok = true;
                else {
                    const unsigned long long fv = tab[x].first;
                    f = ~(uint32_t)fv;
                    ok = ((fv >> 32) == (tag >> 32)) && f < req_off + (uint32_t)i;
                }
            }
            if (ok) { dstar = d0 + k + 1; star_slot = x; star_first = f; }
            else go = false;
        }
    }
    depth_out[i] = dstar;
    if (dstar == 0 || (star_slot & PREOK_BIT)) {
        const uint64_t cand = (dstar == 0) ? avail : (tab[star_slot & ~PREOK_BIT].mask & avail);
        const int cnt = __popcll(cand);
        leader[i] = -1;
        choice[i] = cnt ? nth_set_bit(cand, pick(seed, seq0 + (pos ? (uint64_t)pos[i] : req_off + (uint64_t)i), (uint32_t)cnt)) : -1;
        matched_out[i] = cand;
    } else {
        leader[i] = (int32_t)star_first;
        choice[i] = -1;
    }
}

__global__ void __launch_bounds__(256)
k_select(const Entry *__restrict__ tab, const int32_t *__restrict__ chunk_base, int n, int64_t max_chunks,
         const uint32_t *__restrict__ slots, uint64_t avail, unsigned long long tag, uint64_t seed,
         uint64_t seq0, int32_t *__restrict__ leader, int32_t *__restrict__ choice,
         int32_t *__restrict__ depth_out, uint64_t *__restrict__ matched_out, uint32_t req_off,
         const P2PDyn *__restrict__ dyn, const uint32_t *__restrict__ pos,
         const int32_t *__restrict__ n_dev)
{
    // one thread per request: independent dependent-load chains, 32 per warp in flight.
    // req_off = position of request 0 of this shard in the global (rank-major) order (0 on a
    // single GPU); `first` words, leaders and the tie-break counter are global.  With `dyn`
    // (peer-exchange paths) req_off and seq0 live on the device; with `pos` (partitioned
    // path) pos[i] is the request's global position for the tie-break counter only.
    if (n_dev) n = *n_dev;
    if (dyn) { req_off = dyn->my_off; seq0 = dyn->seq_base; }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        select_request(i, tab, chunk_base, max_chunks, slots, avail, tag, seed, seq0, leader, choice,
                       depth_out, matched_out, req_off, pos);
}

// a5 step 3: followers take their (transitive) leader's endpoint; every
// request ORs its endpoint into the nodes below its match depth -- exactly the
// nodes it was the first to touch, so no other request writes them.
__device__ __forceinline__ void
commit_request(int i, Entry *__restrict__ tab, TrieDev *dev, const int32_t *__restrict__ chunk_base,
               int64_t max_chunks, const uint32_t *__restrict__ slots,
               const int32_t *__restrict__ leader, const int32_t *__restrict__ choice,
               const int32_t *__restrict__ depth, int32_t *__restrict__ endpoint,
               uint64_t *__restrict__ matched_out)
{
    int j = i, l;
    while ((l = leader[j]) >= 0) j = l;          // chains strictly decrease in index
    const int e = choice[j];
    endpoint[i] = e;
    if (j != i) matched_out[i] = (e >= 0) ? (1ULL << e) : 0;
    if (e < 0) return;
    const uint64_t bit = 1ULL << e;
    if (!(dev->root_mask & bit)) atomicOr((unsigned long long *)&dev->root_mask, bit);
    const ReqSpan sp = req_span(chunk_base, i, max_chunks);
    // (fetching the slots eight at a time before the atomics was tried: same time live, slower
    // cold -- the atomics then start later; the plain loop lets them trail the loads)
    for (int d = depth[i]; d < sp.C; d++) {
        const uint32_t x = slots[sp.b0 + d];
        if (x != SLOT_NONE) atomicOr((unsigned long long *)&tab[x & ~PREOK_BIT].mask, bit);
    }
}

__global__ void __launch_bounds__(256)
k_commit(Entry *__restrict__ tab, TrieDev *dev, const int32_t *__restrict__ chunk_base, int n,
         int64_t max_chunks, const uint32_t *__restrict__ slots, const int32_t *__restrict__ leader,
         const int32_t *__restrict__ choice, const int32_t *__restrict__ depth,
         int32_t *__restrict__ endpoint, uint64_t *__restrict__ matched_out,
         const int32_t *__restrict__ n_dev)
{
    if (n_dev) n = *n_dev;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        commit_request(i, tab, dev, chunk_base, max_chunks, slots, leader, choice, depth, endpoint,
                       matched_out);
}

// =========================================================================
// Multi-GPU: request batches sharded across ranks, table replicated per GPU.
// The global batch is the rank-major concatenation of the shards; results must
// equal routing it sequentially.  Per batch every rank
//   1. k_shard_classify  hashes its shard and looks every chunk up in its
//      (identical) replica; chunks whose node is not usable yet become 32-byte
//      records {path key, chunk hash, request index inside the shard, depth};
//   2. the records are all-gathered (NCCL over NVLink; host side);
//   3. k_shard_apply     applies ALL ranks' records to the local replica:
//      find-or-create + atomicMax of the global "first toucher" -> replicas stay
//      identical in content;
//   4. k_select (global positions) on the own shard; leaders / choices / match
//      depths of all ranks are all-gathered;
//   5. k_shard_resolve + k_shard_commit: every rank resolves every request's
//      endpoint and ORs it into the nodes of every record below its owner's
//      match depth -> all replicas receive all inserts.
// =========================================================================
struct __align__(32) ShardRec { uint64_t k1, k2, h; uint32_t req, depth; };

__global__ void __launch_bounds__(WARPS_PER_CTA * 32)
k_shard_classify(const Entry *__restrict__ tab, int cap_log2, const uint64_t *__restrict__ hashes,
                 const int32_t *__restrict__ chunk_base, int n, int64_t max_chunks, uint64_t avail,
                 uint32_t *__restrict__ slots, ShardRec *__restrict__ recs,
                 unsigned int *__restrict__ rec_count)
{
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (i >= n) return;
    const ReqSpan sp = req_span(chunk_base, i, max_chunks);
    Key carry = { KROOT1, KROOT2 };
    for (int blk = 0; blk * 32 < sp.C; blk++) {
        const int d = blk * 32 + lane;
        const bool valid = d < sp.C;
        const uint64_t h = valid ? hashes[sp.b0 + d] : 0;
        const Key k = key_scan_warp(h, carry, lane);
        carry.k1 = __shfl_sync(0xffffffffu, k.k1, 31);
        carry.k2 = __shfl_sync(0xffffffffu, k.k2, 31);
        bool emit = false;
        if (valid) {
            const uint32_t s = table_find(tab, cap_log2, k.k1, k.k2);
            const uint64_t m = (s != SLOT_NONE) ? tab[s].mask : 0;
            if (m & avail) slots[sp.b0 + d] = s | PREOK_BIT;
            else { slots[sp.b0 + d] = SLOT_NONE; emit = true; }
        }
        const unsigned bal = __ballot_sync(0xffffffffu, emit);
        if (bal) {
            unsigned base = 0;
            const int leader = __ffs(bal) - 1;
            if (lane == leader) base = atomicAdd(rec_count, (unsigned)__popc(bal));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (emit) {
                ShardRec r;
                r.k1 = k.k1; r.k2 = k.k2; r.h = h; r.req = (uint32_t)i; r.depth = (uint32_t)d;
                recs[base + __popc(bal & ((1u << lane) - 1))] = r;
            }
        }
    }
}

// Gathered layouts (rank q):
//   meta[2q] = number of records of rank q, meta[2q+1] = number of requests of rank q
//   records  at recs_all[q * stride ...]
//   leaders / choices / match depths at lcd[(3q + 0|1|2) * rstride ...]
// The global position of request i of rank q is req_off(q) + i, req_off = prefix sum of
// the shard sizes (world is small: computed on the fly).
__device__ __forceinline__ uint32_t shard_req_off(const int32_t *meta, int q)
{
    uint32_t o = 0;
    for (int p = 0; p < q; p++) o += (uint32_t)meta[2 * p + 1];
    return o;
}

__global__ void k_shard_apply(Entry *__restrict__ tab, uint64_t *__restrict__ hs, int cap_log2,
                              TrieDev *dev, const ShardRec *__restrict__ recs_all,
                              const int32_t *__restrict__ meta, int world, int64_t stride,
                              int my_rank, const int32_t *__restrict__ chunk_base,
                              unsigned long long tag, uint32_t *__restrict__ slots,
                              uint32_t *__restrict__ rec_slot)
{
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)world * stride) return;
    const int q = (int)(idx / stride);
    if (idx - (int64_t)q * stride >= meta[2 * q]) return;
    const ShardRec r = recs_all[idx];
    bool created;
    const uint32_t s = table_claim(tab, hs, cap_log2, r.k1, r.k2, r.h, &created);
    rec_slot[idx] = s;
    if (s == SLOT_NONE) { atomicOr(&dev->status, B200R_S_TABLE_FULL); return; }
    const unsigned long long mine = tag | (unsigned long long)(~(shard_req_off(meta, q) + r.req));
    unsigned long long cur = 0;
    if (!created) cur = *(volatile unsigned long long *)&tab[s].first;
    if (cur < mine) atomicMax(&tab[s].first, mine);
    if (q == my_rank) slots[chunk_base[r.req] + r.depth] = s;
}

__global__ void k_shard_resolve(TrieDev *dev, const int32_t *__restrict__ lcd,
                                const int32_t *__restrict__ meta, int world, int64_t rstride,
                                int32_t *__restrict__ endpoint_all)
{
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)world * rstride) return;
    int q = (int)(idx / rstride);
    int64_t i = idx - (int64_t)q * rstride;
    if (i >= meta[2 * q + 1]) return;
    int l;
    while ((l = lcd[(3 * (int64_t)q) * rstride + i]) >= 0) {   // leader = global position
        uint32_t pos = (uint32_t)l, o = 0;
        q = 0;
        while (q + 1 < world && pos >= o + (uint32_t)meta[2 * q + 1]) { o += (uint32_t)meta[2 * q + 1]; q++; }
        i = pos - o;
    }
    const int e = lcd[(3 * (int64_t)q + 1) * rstride + i];
    endpoint_all[idx] = e;
    if (e >= 0) {
        const uint64_t bit = 1ULL << e;
        if (!(dev->root_mask & bit)) atomicOr((unsigned long long *)&dev->root_mask, bit);
    }
}

__global__ void k_shard_commit(Entry *__restrict__ tab, const ShardRec *__restrict__ recs_all,
                               const int32_t *__restrict__ meta, int world, int64_t stride,
                               const uint32_t *__restrict__ rec_slot,
                               const int32_t *__restrict__ endpoint_all,
                               const int32_t *__restrict__ lcd, int64_t rstride)
{
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)world * stride) return;
    const int q = (int)(idx / stride);
    if (idx - (int64_t)q * stride >= meta[2 * q]) return;
    const uint32_t s = rec_slot[idx];
    if (s == SLOT_NONE) return;
    const uint32_t req = recs_all[idx].req, d = recs_all[idx].depth;
    const int e = endpoint_all[(int64_t)q * rstride + req];
    if (e >= 0 && (int)d >= lcd[(3 * (int64_t)q + 2) * rstride + req])
        atomicOr((unsigned long long *)&tab[s].mask, 1ULL << e);
}

__global__ void k_shard_finish(const int32_t *__restrict__ endpoint_all,
                               const int32_t *__restrict__ lcd, int n, int my_rank, int64_t rstride,
                               int32_t *__restrict__ endpoint, int32_t *__restrict__ depth,
                               uint64_t *__restrict__ matched)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int e = endpoint_all[(int64_t)my_rank * rstride + i];
    endpoint[i] = e;
    depth[i] = lcd[(3 * (int64_t)my_rank + 2) * rstride + i];
    if (lcd[(3 * (int64_t)my_rank) * rstride + i] >= 0) matched[i] = (e >= 0) ? (1ULL << e) : 0;
}

// =========================================================================
// Multi-GPU exchange over NVLink peer memory (no NCCL, no host round trip).
// Every rank owns one "symmetric" buffer that all peers can address:
//   P2PFlags                      arrival flags, one set per batch parity
//   recs  [2][world][cap]         32-byte records, region q written by rank q
//   lcd   [2][world][3][rcap]     leaders / choices / depths, region q by rank q
// A rank PUSHES its records (and later its leaders/choices/depths) into the
// region it owns in every peer's buffer with coalesced 16-byte stores, then
// publishes (epoch << 32 | count) flags; consumers spin on their own flags.  Two
// buffer sets (batch parity) suffice: a rank can run at most one batch ahead of
// a peer, because its next classify needs that peer's leaders of this batch.
// =========================================================================
struct P2PFlags {
    unsigned long long rec_cnt[2][P2P_MAX_WORLD], rec_n[2][P2P_MAX_WORLD], lcd[2][P2P_MAX_WORLD];
};
constexpr size_t P2P_FLAGS_BYTES = 4096;
static_assert(sizeof(P2PFlags) <= P2P_FLAGS_BYTES, "flag block");

__global__ void k_p2p_push(const uint4 *__restrict__ src, const unsigned int *__restrict__ count_ptr,
                           long long fixed_items, int units_per_item,
                           const unsigned long long *__restrict__ peer_bases, int world,
                           size_t dst_off)
{
    const long long items = count_ptr ? (long long)*count_ptr : fixed_items;
    const long long units = items * units_per_item;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < units;
         i += (long long)gridDim.x * blockDim.x) {
        const uint4 v = src[i];
        for (int p = 0; p < world; p++)
            ((uint4 *)(uintptr_t)(peer_bases[p] + dst_off))[i] = v;
    }
    __threadfence_system();
}

__global__ void k_p2p_publish(const unsigned long long *__restrict__ peer_bases, int world,
                              int my_rank, int set, unsigned long long epoch, int kind,
                              const unsigned int *__restrict__ count_ptr, unsigned int n)
{
    const int p = threadIdx.x;
    if (p >= world) return;
    P2PFlags *f = (P2PFlags *)(uintptr_t)peer_bases[p];
    __threadfence_system();
    if (kind == 0) {
        *(volatile unsigned long long *)&f->rec_n[set][my_rank] = (epoch << 32) | n;
        *(volatile unsigned long long *)&f->rec_cnt[set][my_rank] = (epoch << 32) | *count_ptr;
    } else {
        *(volatile unsigned long long *)&f->lcd[set][my_rank] = (epoch << 32) | n;
    }
    __threadfence_system();
}

__device__ __forceinline__ bool p2p_spin(const volatile unsigned long long *flag,
                                         unsigned long long epoch, unsigned long long *val)
{
    const long long t0 = clock64();
    for (;;) {
        const unsigned long long v = *flag;
        if ((v >> 32) == epoch) { *val = v; return true; }
        if (clock64() - t0 > 6000000000LL) return false;     // ~3 s: a peer is gone
        __nanosleep(200);
    }
}

__global__ void k_p2p_wait_meta(const P2PFlags *__restrict__ mine, int set, unsigned long long epoch,
                                int world, int my_rank, P2PDyn *dyn, TrieDev *dev)
{
    const int q = threadIdx.x;
    if (q < world) {
        unsigned long long c = 0, m = 0;
        const bool ok = p2p_spin(&mine->rec_cnt[set][q], epoch, &c) &&
                        p2p_spin(&mine->rec_n[set][q], epoch, &m);
        if (!ok) atomicOr(&dev->status, B200R_S_PEER_TIMEOUT);
        dyn->meta[2 * q] = ok ? (int32_t)(uint32_t)c : 0;
        dyn->meta[2 * q + 1] = ok ? (int32_t)(uint32_t)m : 0;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t off = 0, tot = 0;
        for (int r = 0; r < world; r++) {
            if (r == my_rank) off = tot;
            tot += (uint32_t)dyn->meta[2 * r + 1];
        }
        dyn->my_off = off;
        dyn->total_n = tot;
    }
    __threadfence_system();
}

__global__ void k_p2p_wait_lcd(const P2PFlags *__restrict__ mine, int set, unsigned long long epoch,
                               int world, TrieDev *dev)
{
    const int q = threadIdx.x;
    if (q < world) {
        unsigned long long v;
        if (!p2p_spin(&mine->lcd[set][q], epoch, &v)) atomicOr(&dev->status, B200R_S_PEER_TIMEOUT);
    }
    __threadfence_system();
}

__global__ void k_p2p_advance(P2PDyn *dyn) { dyn->seq_base += dyn->total_n; }

// =========================================================================
// Multi-GPU, PARTITIONED table (the scalable variant; SURVEY.md 8e fallback).
// A request's whole path hangs under its first chunk, so requests with
// different first-chunk hashes never meet in the trie.  The table is therefore
// partitioned by owner(first chunk hash) = mix(h0) % world and nothing is
// replicated: per batch every rank
//   send    hashes its own shard (as before) and pushes, per owner, the chunk
//           hashes + {local index, chunk count} of the requests that owner
//           holds, order preserved, straight into the owner's buffer (NVLink
//           peer stores, ~6 % of the prompt bytes), then publishes flags;
//   route   waits for all sources, concatenates what it received in rank
//           order (= global order of the requests it owns), runs the ordinary
//           claim / select / commit on that batch against ITS partition, and
//           pushes {endpoint, depth, candidates} back to the sources;
//   finish  waits for all owners and scatters the results to request order.
// Sequential equivalence holds per partition and partitions are independent;
// the tie-break counter uses the request's position in the global order.
// Per-GPU work no longer grows with the number of GPUs.
// =========================================================================
struct PartFlags {
    unsigned long long in_req[2][P2P_MAX_WORLD], in_chunk[2][P2P_MAX_WORLD], in_n[2][P2P_MAX_WORLD],
        res[2][P2P_MAX_WORLD];
};
static_assert(sizeof(PartFlags) <= 4096, "flag block");
struct PartRes { int32_t ep, depth; uint64_t matched; };
struct PartDyn {
    P2PDyn base;                              // seq_base, total_n (my_off stays 0)
    int32_t n_owned, c_owned;
    uint32_t reqbase[P2P_MAX_WORLD + 1], chunkbase[P2P_MAX_WORLD + 1], off[P2P_MAX_WORLD + 1];
};

__device__ __forceinline__ int part_owner(const uint64_t *hashes, int b, int C, int world)
{
    if (C == 0) return 0;                     // empty prompt: touches only a root
    uint64_t x = hashes[b] * 0x9E3779B97F4A7C15ULL;
    return (int)((x >> 33) % (uint64_t)world);
}

constexpr int PART_BLOCK = 256;

// per block of 256 requests: how many requests / chunks go to each owner
__global__ void __launch_bounds__(PART_BLOCK)
k_part_count(const int32_t *__restrict__ chunk_base, const uint64_t *__restrict__ hashes, int n,
             int world, int32_t *__restrict__ blk_cnt)
{
    __shared__ int s_cnt[2 * P2P_MAX_WORLD];
    for (int k = threadIdx.x; k < 2 * world; k += PART_BLOCK) s_cnt[k] = 0;
    __syncthreads();
    const int i = blockIdx.x * PART_BLOCK + threadIdx.x;
    if (i < n) {
        const int b = chunk_base[i], C = chunk_base[i + 1] - b;
        const int o = part_owner(hashes, b, C, world);
        atomicAdd(&s_cnt[2 * o], 1);
        atomicAdd(&s_cnt[2 * o + 1], C);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < 2 * world; k += PART_BLOCK)
        blk_cnt[(size_t)blockIdx.x * 2 * world + k] = s_cnt[k];
}

// exclusive scan over the blocks, per owner and kind (one warp each, 32 blocks per step)
__global__ void k_part_scan(const int32_t *__restrict__ blk_cnt, int nblk, int world,
                            int32_t *__restrict__ blk_off, int32_t *__restrict__ totals)
{
    const int lane = threadIdx.x & 31;
    const int nwarps = blockDim.x >> 5;
    for (int k = threadIdx.x >> 5; k < 2 * world; k += nwarps) {
        int run = 0;
        for (int b0 = 0; b0 < nblk; b0 += 32) {
            const int b = b0 + lane;
            const int v = b < nblk ? blk_cnt[(size_t)b * 2 * world + k] : 0;
            int inc = v;
            for (int sft = 1; sft < 32; sft <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, inc, sft);
                if (lane >= sft) inc += t;
            }
            if (b < nblk) blk_off[(size_t)b * 2 * world + k] = run + inc - v;
            run += __shfl_sync(0xffffffffu, inc, 31);
        }
        if (lane == 0) totals[k] = run;
    }
    // starts of the per-owner segments in the local staging buffers (even, so that every
    // segment can be copied with 16-byte vectors): totals[2*world + 2*o] requests, [+1] chunks
    __syncthreads();
    if (threadIdx.x == 0) {
        int r = 0, c = 0;
        for (int o = 0; o < world; o++) {
            totals[2 * world + 2 * o] = r; totals[2 * world + 2 * o + 1] = c;
            r += (totals[2 * o] + 1) & ~1; c += (totals[2 * o + 1] + 1) & ~1;
        }
    }
}

// order-preserving compaction per owner: where does every request (and its run of chunk
// hashes) go in its owner's buffer?  The {local index, chunk count} words and the hashes
// are first compacted into LOCAL staging buffers, one contiguous segment per owner, and then
// shipped by k_part_bulk with full-line 16-byte vector stores: NVLink moves the same bytes
// several times faster in 128-byte lines than as 8-byte words or ragged ~170-byte runs
// (measured at 2 GPUs: 47 us per shard for direct scattered peer stores).
__global__ void __launch_bounds__(PART_BLOCK)
k_part_push(const int32_t *__restrict__ chunk_base, const uint64_t *__restrict__ hashes, int n,
            int world, const int32_t *__restrict__ blk_off, const int32_t *__restrict__ totals,
            unsigned long long *__restrict__ info_stage, uint32_t *__restrict__ dest_slot,
            uint32_t *__restrict__ req_dst)
{
    __shared__ int s_w[PART_BLOCK / 32][2 * P2P_MAX_WORLD];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int i = blockIdx.x * PART_BLOCK + threadIdx.x;
    int b = 0, C = 0, o = -1;
    if (i < n) {
        b = chunk_base[i];
        C = chunk_base[i + 1] - b;
        o = part_owner(hashes, b, C, world);
    }
    // rank of this request / its chunks among the warp's earlier requests of the same owner
    int rreq = 0, rchunk = 0;
    for (int d = 0; d < world; d++) {
        const unsigned m = __ballot_sync(0xffffffffu, o == d);
        int v = (o == d) ? C : 0, inc = v;
        for (int sft = 1; sft < 32; sft <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, sft);
            if (lane >= sft) inc += t;
        }
        if (o == d) { rreq = __popc(m & ((1u << lane) - 1)); rchunk = inc - v; }
        if (lane == 31) { s_w[w][2 * d] = __popc(m); s_w[w][2 * d + 1] = inc; }
    }
    __syncthreads();
    if (o >= 0) {
        int breq = blk_off[(size_t)blockIdx.x * 2 * world + 2 * o];
        int bchunk = blk_off[(size_t)blockIdx.x * 2 * world + 2 * o + 1];
        for (int k = 0; k < w; k++) { breq += s_w[k][2 * o]; bchunk += s_w[k][2 * o + 1]; }
        rreq += breq;
        rchunk += bchunk;
        dest_slot[i] = ((uint32_t)o << 24) | (uint32_t)rreq;
        req_dst[i] = (uint32_t)(totals[2 * world + 2 * o + 1] + rchunk);
        info_stage[totals[2 * world + 2 * o] + rreq] =
            (unsigned long long)(uint32_t)i | ((unsigned long long)(uint32_t)C << 32);
    }
}

// chunk hashes -> local staging, one warp per request (coalesced both ways)
__global__ void __launch_bounds__(WARPS_PER_CTA * 32)
k_part_copy(const int32_t *__restrict__ chunk_base, const uint64_t *__restrict__ hashes, int n,
            const uint32_t *__restrict__ req_dst, uint64_t *__restrict__ hash_stage)
{
    const int lane = threadIdx.x & 31;
    const int nw = (gridDim.x * blockDim.x) >> 5;
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += nw) {
        const int b = chunk_base[i], C = chunk_base[i + 1] - b;
        uint64_t *dst = hash_stage + req_dst[i];
        for (int c = lane; c < C; c += 32) dst[c] = hashes[b + c];
    }
}

// staging segments -> the owners' buffers: 16-byte vectors, whole lines.  Persistent grid, one
// system-scope fence per CTA at the end (MEMBAR.SYS costs microseconds: not once per wave).
__global__ void __launch_bounds__(256)
k_part_bulk(const uint64_t *__restrict__ hash_stage, const unsigned long long *__restrict__ info_stage,
            const int32_t *__restrict__ totals, int world, int my_rank,
            const unsigned long long *__restrict__ peer_bases, size_t hash_off, size_t info_off,
            int64_t cap, int64_t rcap)
{
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
    for (int k = 0; k < world; k++) {
        const int o = (my_rank + 1 + k) % world;       // start with the next rank: spread the links
        const int nreq = totals[2 * o], nchunk = totals[2 * o + 1];
        const uint4 *hs = (const uint4 *)(hash_stage + totals[2 * world + 2 * o + 1]);
        uint4 *hd = (uint4 *)((uint64_t *)(uintptr_t)(peer_bases[o] + hash_off) + (size_t)my_rank * cap);
        for (int64_t x = tid; x < (nchunk + 1) / 2; x += nthr) hd[x] = hs[x];
        const uint4 *is = (const uint4 *)(info_stage + totals[2 * world + 2 * o]);
        uint4 *id = (uint4 *)((unsigned long long *)(uintptr_t)(peer_bases[o] + info_off) + (size_t)my_rank * rcap);
        for (int64_t x = tid; x < (nreq + 1) / 2; x += nthr) id[x] = is[x];
    }
    __syncthreads();
    if (threadIdx.x == 0) __threadfence_system();
}

__global__ void k_part_publish_in(const unsigned long long *__restrict__ peer_bases, int world,
                                  int my_rank, int set, unsigned long long epoch,
                                  const int32_t *__restrict__ totals, unsigned int n)
{
    const int d = threadIdx.x;
    if (d >= world) return;
    PartFlags *f = (PartFlags *)(uintptr_t)peer_bases[d];
    __threadfence_system();
    *(volatile unsigned long long *)&f->in_n[set][my_rank] = (epoch << 32) | n;
    *(volatile unsigned long long *)&f->in_chunk[set][my_rank] = (epoch << 32) | (uint32_t)totals[2 * d + 1];
    *(volatile unsigned long long *)&f->in_req[set][my_rank] = (epoch << 32) | (uint32_t)totals[2 * d];
    __threadfence_system();
}

__global__ void k_part_wait_meta(const PartFlags *__restrict__ mine, int set, unsigned long long epoch,
                                 int world, PartDyn *dyn, TrieDev *dev)
{
    __shared__ uint32_t s_req[P2P_MAX_WORLD], s_chunk[P2P_MAX_WORLD], s_n[P2P_MAX_WORLD];
    const int q = threadIdx.x;
    if (q < world) {
        unsigned long long a = 0, b = 0, c = 0;
        const bool ok = p2p_spin(&mine->in_req[set][q], epoch, &a) &&
                        p2p_spin(&mine->in_chunk[set][q], epoch, &b) &&
                        p2p_spin(&mine->in_n[set][q], epoch, &c);
        if (!ok) atomicOr(&dev->status, B200R_S_PEER_TIMEOUT);
        s_req[q] = ok ? (uint32_t)a : 0; s_chunk[q] = ok ? (uint32_t)b : 0; s_n[q] = ok ? (uint32_t)c : 0;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t r = 0, c = 0, o = 0;
        for (int k = 0; k < world; k++) {
            dyn->reqbase[k] = r; dyn->chunkbase[k] = c; dyn->off[k] = o;
            r += s_req[k]; c += s_chunk[k]; o += s_n[k];
        }
        dyn->reqbase[world] = r; dyn->chunkbase[world] = c; dyn->off[world] = o;
        dyn->n_owned = (int32_t)r; dyn->c_owned = (int32_t)c;
        dyn->base.my_off = 0; dyn->base.total_n = o;
    }
    __threadfence_system();
}

// concatenate what the sources sent, in rank order
__global__ void k_part_gather(const uint8_t *__restrict__ my_base, size_t hash_off, size_t info_off,
                              int64_t cap, int64_t rcap, int world, const PartDyn *__restrict__ dyn,
                              int32_t *__restrict__ c_owned, uint32_t *__restrict__ pos_owned,
                              uint64_t *__restrict__ hashes_owned)
{
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t step = (int64_t)gridDim.x * blockDim.x;
    const unsigned long long *info = (const unsigned long long *)(my_base + info_off);
    const uint64_t *hin = (const uint64_t *)(my_base + hash_off);
    for (int64_t j = tid; j < dyn->n_owned; j += step) {
        int q = 0;
        while (q + 1 < world && j >= dyn->reqbase[q + 1]) q++;
        const unsigned long long v = info[(size_t)q * rcap + (j - dyn->reqbase[q])];
        c_owned[j] = (int32_t)(v >> 32);
        pos_owned[j] = dyn->off[q] + (uint32_t)v;
    }
    for (int64_t x = tid; x < dyn->c_owned; x += step) {
        int q = 0;
        while (q + 1 < world && x >= dyn->chunkbase[q + 1]) q++;
        hashes_owned[x] = hin[(size_t)q * cap + (x - dyn->chunkbase[q])];
    }
}

__global__ void k_part_push_results(const PartDyn *__restrict__ dyn, int world, int my_rank,
                                    const int32_t *__restrict__ ep, const int32_t *__restrict__ depth,
                                    const uint64_t *__restrict__ matched,
                                    const unsigned long long *__restrict__ peer_bases,
                                    size_t res_off, int64_t rcap)
{
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < dyn->n_owned;
         j += (int64_t)gridDim.x * blockDim.x) {
        int q = 0;
        while (q + 1 < world && j >= dyn->reqbase[q + 1]) q++;
        PartRes r;
        r.ep = ep[j]; r.depth = depth[j]; r.matched = matched[j];
        ((PartRes *)(uintptr_t)(peer_bases[q] + res_off))[(size_t)my_rank * rcap + (j - dyn->reqbase[q])] = r;
    }
    __syncthreads();
    if (threadIdx.x == 0) __threadfence_system();
}

__global__ void k_part_publish_res(const unsigned long long *__restrict__ peer_bases, int world,
                                   int my_rank, int set, unsigned long long epoch,
                                   const PartDyn *__restrict__ dyn)
{
    const int q = threadIdx.x;
    if (q >= world) return;
    PartFlags *f = (PartFlags *)(uintptr_t)peer_bases[q];
    __threadfence_system();
    *(volatile unsigned long long *)&f->res[set][my_rank] =
        (epoch << 32) | (dyn->reqbase[q + 1] - dyn->reqbase[q]);
    __threadfence_system();
}

__global__ void k_part_wait_res(const PartFlags *__restrict__ mine, int set, unsigned long long epoch,
                                int world, TrieDev *dev, PartDyn *dyn)
{
    const int q = threadIdx.x;
    if (q < world) {
        unsigned long long v;
        if (!p2p_spin(&mine->res[set][q], epoch, &v)) atomicOr(&dev->status, B200R_S_PEER_TIMEOUT);
    }
    __syncthreads();
    if (threadIdx.x == 0) dyn->base.seq_base += dyn->base.total_n;   // next batch's request 0
    __threadfence_system();
}

__global__ void k_part_scatter(const uint32_t *__restrict__ dest_slot, const uint8_t *__restrict__ my_base,
                               size_t res_off, int64_t rcap, int n, int32_t *__restrict__ endpoint,
                               int32_t *__restrict__ depth, uint64_t *__restrict__ matched)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t ds = dest_slot[i];
    const PartRes r = ((const PartRes *)(my_base + res_off))[(size_t)(ds >> 24) * rcap + (ds & 0xFFFFFFu)];
    endpoint[i] = r.ep; depth[i] = r.depth; matched[i] = r.matched;
}

// ------------------------------------------------------------ dump / restore
__global__ void k_count_nodes(const Entry *__restrict__ tab, int64_t cap, TrieDev *dev)
{
    unsigned long long mine = 0;
    for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < cap;
         s += (int64_t)gridDim.x * blockDim.x) {
        uint64_t a, b;
        ld_key(tab + s, a, b);
        mine += (a | b) ? 1 : 0;
    }
    for (int o = 16; o; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(&dev->num_nodes, mine);
}

__global__ void k_dump(const Entry *__restrict__ tab, const uint64_t *__restrict__ hs, int64_t cap, TrieDev *dev,
                       uint64_t *__restrict__ out, int64_t max_nodes)
{
    for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < cap;
         s += (int64_t)gridDim.x * blockDim.x) {
        const Entry e = tab[s];
        if (e.k1 | e.k2) {
            unsigned long long idx = atomicAdd(&dev->dump_count, 1ULL);
            if ((int64_t)idx < max_nodes) {
                out[4 * idx + 0] = e.k1;
                out[4 * idx + 1] = e.k2;
                out[4 * idx + 2] = e.h;
                out[4 * idx + 3] = e.mask;
            }
        }
    }
}

__global__ void k_apply_remote(Entry *__restrict__ tab, uint64_t *__restrict__ hs, int cap_log2, TrieDev *dev,
                               const uint64_t *__restrict__ ent, int64_t count, uint64_t root_or)
{
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx == 0 && root_or) atomicOr((unsigned long long *)&dev->root_mask, root_or);
    bool created = false;
    if (idx < count) {
        const uint64_t k1 = ent[4 * idx], k2 = ent[4 * idx + 1], h = ent[4 * idx + 2],
                       m = ent[4 * idx + 3];
        uint32_t s = table_claim(tab, hs, cap_log2, k1, k2, h, &created);
        if (s == SLOT_NONE) atomicOr(&dev->status, B200R_S_TABLE_FULL);
        else if (m) atomicOr((unsigned long long *)&tab[s].mask, m);
    }
}

}  // namespace b200r

// =========================================================================
// C ABI
// =========================================================================
using namespace b200r;

struct b200r_trie {
    int device;
    int cap_log2;
    int chunk_chars;
    Entry *tab;
    uint64_t *hs;
    TrieDev *dev;
    uint32_t epoch;
};

static inline size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }

struct Plan {           // carve-up of the caller's scratch buffer
    size_t off_tile_first, off_block_sums, off_nchars, off_chunk_off, off_chunk_base, off_hashes,
        off_slots, off_leader, off_choice, off_ticket, off_lb_flag, off_lb_part, off_lb_final, off_chunk_req, total;
};

static Plan make_plan(int64_t n, int64_t max_chunks, uint32_t flags, bool routing)
{
    Plan p;
    size_t o = 0;
    const int64_t ntiles = (max_chunks + 31) / 32 + 1;
    const int64_t nblk = (n + PREP_BLOCK - 1) / PREP_BLOCK + 1;
    const bool ascii = flags & B200R_F_ALL_ASCII;
    p.off_tile_first = o; o += align_up(ntiles * 4);
    p.off_ticket = o;     o += 256;
    p.off_lb_flag = o;    o += align_up(ntiles * 4);
    p.off_lb_part = o;    o += routing ? align_up(ntiles * 32) : 0;
    p.off_lb_final = o;   o += routing ? align_up(ntiles * 32) : 0;
    p.off_block_sums = o; o += align_up(nblk * 4);
    p.off_nchars = o;     o += ascii ? 0 : align_up((n + 1) * 4);
    p.off_chunk_off = o;  o += ascii ? 0 : align_up((max_chunks + 1) * 4);
    p.off_chunk_base = o; o += routing ? align_up((n + 1) * 4) : 0;
    p.off_hashes = o;     o += routing ? align_up((max_chunks + 1) * 8) : 0;
    p.off_slots = o;      o += routing ? align_up((max_chunks + 1) * 4) : 0;
    p.off_leader = o;     o += routing ? align_up((n + 1) * 4) : 0;
    p.off_choice = o;     o += routing ? align_up((n + 1) * 4) : 0;
    p.off_chunk_req = o;  o += routing ? align_up((max_chunks + 1) * 4) : 0;
    p.total = o + 256;
    return p;
}

static int check_prompt_args(const void *bytes, const void *starts, const void *lens, int64_t n,
                             int64_t max_chunks, int32_t cc)
{
    if (n < 0 || max_chunks < 0 || cc <= 0 || n > 0x7fffffff - 2048 || max_chunks > 0x7fffffff - 2048) {
        set_error("invalid sizes: n=%lld max_chunks=%lld chunk_chars=%d", (long long)n,
                  (long long)max_chunks, cc);
        return B200R_E_INVALID;
    }
    if (n > 0 && (!bytes || !starts || !lens)) {
        set_error("null prompt pointer");
        return B200R_E_INVALID;
    }
    return B200R_OK;
}

// prep + (utf8) + hash; shared by b200r_xxh64_chunks and b200r_prefix_route_batch
static int launch_hash(const uint8_t *d_bytes, const int64_t *d_starts, const int32_t *d_lens,
                       const int32_t *d_nchars, int64_t n, int32_t cc, uint32_t flags,
                       int32_t *d_chunk_base, uint64_t *d_hashes, int64_t max_chunks,
                       uint8_t *scratch, const Plan &p, TrieDev *dev, cudaStream_t st,
                       bool prep_only = false, int32_t *chunk_req = nullptr)
{
    int32_t *tile_first = (int32_t *)(scratch + p.off_tile_first);
    int32_t *block_sums = (int32_t *)(scratch + p.off_block_sums);
    int32_t *chunk_off = nullptr;
    const bool ascii = flags & B200R_F_ALL_ASCII;
    const int nwarp_blocks = (int)((n + WARPS_PER_CTA - 1) / WARPS_PER_CTA);
    if (ascii) {
        d_nchars = nullptr;                     // nchars == lens by contract
    } else {
        chunk_off = (int32_t *)(scratch + p.off_chunk_off);
        if (!d_nchars) {
            int32_t *tmp = (int32_t *)(scratch + p.off_nchars);
            { ProfScope ps__(PK_UTF8, st); g_launches += 1;
            k_count_chars<<<nwarp_blocks, WARPS_PER_CTA * 32, 0, st>>>(d_bytes, d_starts, d_lens,
                                                                      (int)n, tmp);
            B200R_LAUNCH_CHECK("k_count_chars");
            }
            d_nchars = tmp;
        }
    }
    const int nblk = (int)((n + PREP_BLOCK - 1) / PREP_BLOCK);
    { ProfScope ps__(PK_PREP, st); g_launches += 1;
    k_prep_sums<<<nblk, PREP_BLOCK, 0, st>>>(d_lens, d_nchars, (int)n, cc, block_sums, nullptr,
                                             (unsigned int *)(scratch + p.off_ticket));
    B200R_LAUNCH_CHECK("k_prep_sums");
    }
    { ProfScope ps__(PK_PREP, st); g_launches += 1;
    k_prep_scan<<<nblk, PREP_BLOCK, 0, st>>>(d_lens, d_nchars, (int)n, cc, block_sums, d_chunk_base,
                                             tile_first, max_chunks, dev, d_bytes, d_starts, d_hashes, nullptr,
                                             (unsigned int *)(scratch + p.off_lb_flag));
    B200R_LAUNCH_CHECK("k_prep_scan");
    }
    if (!ascii) {
        { ProfScope ps__(PK_UTF8, st); g_launches += 1;
        k_utf8_bounds<<<nwarp_blocks, WARPS_PER_CTA * 32, 0, st>>>(
            d_bytes, d_starts, d_lens, d_nchars, d_chunk_base, (int)n, cc, max_chunks, chunk_off);
        B200R_LAUNCH_CHECK("k_utf8_bounds");
        }
    }
    const int ntiles = (int)((max_chunks + HASH_WARPS * WTILE - 1) / (HASH_WARPS * WTILE));
    if (ntiles > 0 && !prep_only) {
        { ProfScope ps__(PK_HASH, st); g_launches += 1;
        FuseArgs fa0 = FuseArgs();
        fa0.chunk_req = chunk_req;          // chunk -> request map for the tile claim kernel
        k_hash_tiles<0><<<ntiles, HASH_WARPS * 32, 0, st>>>(
            d_bytes, d_starts, d_lens, d_nchars, d_chunk_base, tile_first, chunk_off, (int)n, cc,
            max_chunks, d_hashes, fa0);
        B200R_LAUNCH_CHECK("k_hash_chunks");
        }
    }
    return B200R_OK;
}

extern "C" {

int b200r_abi_version(void) { return B200R_ABI_VERSION; }
const char *b200r_last_error(void) { return g_err; }
uint32_t b200r_pick(uint64_t seed, uint64_t seq, uint32_t n) { return pick(seed, seq, n); }

int b200r_set_claim_by_tiles(int on)
{
    g_claim_by_tiles = on != 0;
    return B200R_OK;
}

int b200r_prof_enable(int on)
{
    g_prof_on = on != 0;
    return B200R_OK;
}

long long b200r_launch_count(void) { return g_launches; }

int b200r_prof_read(float *ms_out, long long *count_out, int nslots)
{
    float ms[PK_COUNT] = { 0 };
    long long cnt[PK_COUNT] = { 0 };
    for (auto &sp : g_spans) {
        B200R_CUDA(cudaEventSynchronize(sp.b));
        float t = 0;
        B200R_CUDA(cudaEventElapsedTime(&t, sp.a, sp.b));
        ms[sp.kid] += t; cnt[sp.kid]++;
        cudaEventDestroy(sp.a); cudaEventDestroy(sp.b);
    }
    g_spans.clear();
    for (int k = 0; k < nslots && k < PK_COUNT; k++) {
        if (ms_out) ms_out[k] = ms[k];
        if (count_out) count_out[k] = cnt[k];
    }
    return B200R_OK;
}

void b200r_trie_key_constants(uint64_t out[6])
{
    out[0] = KA1; out[1] = KA2; out[2] = KROOT1; out[3] = KROOT2;
    out[4] = cinv(KA1); out[5] = cinv(KA2);
}

size_t b200r_xxh64_chunks_scratch_bytes(int64_t n, int64_t max_chunks)
{
    return make_plan(n, max_chunks, 0, false).total;
}

int b200r_xxh64_chunks(const uint8_t *d_bytes, const int64_t *d_starts, const int32_t *d_lens,
                       const int32_t *d_nchars, int64_t n, int32_t chunk_chars, uint32_t flags,
                       int32_t *d_chunk_base, uint64_t *d_hashes, int64_t max_chunks,
                       void *d_scratch, size_t scratch_bytes, void *stream)
{
    int rc = check_prompt_args(d_bytes, d_starts, d_lens, n, max_chunks, chunk_chars);
    if (rc) return rc;
    if (!d_chunk_base || (max_chunks > 0 && !d_hashes)) { set_error("null output pointer"); return B200R_E_INVALID; }
    const Plan p = make_plan(n, max_chunks, flags, false);
    if (!d_scratch || scratch_bytes < p.total) {
        set_error("scratch too small: need %zu bytes, got %zu", p.total, scratch_bytes);
        return B200R_E_SCRATCH;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) {
        B200R_CUDA(cudaMemsetAsync(d_chunk_base, 0, 4, st));
        return B200R_OK;
    }
    return launch_hash(d_bytes, d_starts, d_lens, d_nchars, n, chunk_chars, flags, d_chunk_base,
                       d_hashes, max_chunks, (uint8_t *)d_scratch, p, nullptr, st);
}

// ---------------------------------------------------------------------------
// Chunk hashing for callers that cannot allocate device memory.
//
// Some users of this library only ever hold the input and output buffers their
// host framework hands them: a router plugin living inside a serving process
// whose engine has already reserved (nearly) all of HBM, so that a cudaMalloc
// for a few MB of work space fails; or a request-path thread that must not call
// cudaMalloc/cudaFree because those calls synchronise the whole device.  For
// them the work space of the hashing pipeline (tile map, block sums, code-point
// counts, UTF-8 chunk offsets) lives in the fixed array below, which is part of
// the library image: every GPU that loads libb200router.so carries these
// B200R_NOALLOC_SCRATCH_BYTES until the process exits.
//
// In host code `g_noalloc_scratch` is only a handle for the symbol, not a device
// address, so the address is looked up with cudaGetSymbolAddress -- once per
// device, with that device current -- and handed to the kernels like any other
// scratch pointer.
static __device__ __align__(256) unsigned char g_noalloc_scratch[B200R_NOALLOC_SCRATCH_BYTES];

static std::mutex g_noalloc_mu;
static uint8_t *g_noalloc_ptr[64] = { nullptr };   // resolved address, per device ordinal

static int noalloc_scratch_for_current_device(uint8_t **out)
{
    int devid = -1;
    B200R_CUDA(cudaGetDevice(&devid));
    if (devid < 0 || devid >= 64) {
        set_error("b200r_xxh64_chunks_noalloc: device ordinal %d out of range", devid);
        return B200R_E_INVALID;
    }
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[devid]) {
        void *ptr = nullptr;
        B200R_CUDA(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
        if (!ptr) { set_error("cudaGetSymbolAddress returned a null address"); return B200R_E_CUDA; }
        g_noalloc_ptr[devid] = (uint8_t *)ptr;
    }
    *out = g_noalloc_ptr[devid];
    return B200R_OK;
}

int b200r_xxh64_chunks_noalloc(const uint8_t *d_bytes, const int64_t *d_starts,
                               const int32_t *d_lens, const int32_t *d_nchars, int64_t n,
                               int32_t chunk_chars, uint32_t flags, int32_t *d_chunk_base,
                               uint64_t *d_hashes, int64_t max_chunks, void *stream)
{
    int rc = check_prompt_args(d_bytes, d_starts, d_lens, n, max_chunks, chunk_chars);
    if (rc) return rc;
    if (!d_chunk_base || (max_chunks > 0 && !d_hashes)) { set_error("null output pointer"); return B200R_E_INVALID; }
    const Plan p = make_plan(n, max_chunks, flags, false);
    if (p.total > (size_t)B200R_NOALLOC_SCRATCH_BYTES) {
        set_error("b200r_xxh64_chunks_noalloc: a batch of n=%lld requests / max_chunks=%lld chunks "
                  "needs %zu bytes of work space, the built-in area holds %d; split the batch or "
                  "use b200r_xxh64_chunks with a caller-provided scratch buffer",
                  (long long)n, (long long)max_chunks, p.total, B200R_NOALLOC_SCRATCH_BYTES);
        return B200R_E_SCRATCH;
    }
    uint8_t *scratch = nullptr;
    rc = noalloc_scratch_for_current_device(&scratch);   // first call on a device resolves the symbol
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) {
        B200R_CUDA(cudaMemsetAsync(d_chunk_base, 0, 4, st));
        return B200R_OK;
    }
    // Nothing in the work space has to start at zero: every word the kernels
    // read in this call was written earlier in this call (tile map and block
    // sums by the prep kernels, counts/offsets by the UTF-8 kernels), so stale
    // contents of the previous call are never observed.
    return launch_hash(d_bytes, d_starts, d_lens, d_nchars, n, chunk_chars, flags, d_chunk_base,
                       d_hashes, max_chunks, scratch, p, nullptr, st);
}

int b200r_trie_create(int32_t cap_log2, int32_t chunk_chars, b200r_trie **out)
{
    if (!out || cap_log2 < 4 || cap_log2 > 30 || chunk_chars <= 0) {
        set_error("b200r_trie_create: cap_log2 must be 4..30, chunk_chars > 0");
        return B200R_E_INVALID;
    }
    b200r_trie *t = new b200r_trie();
    B200R_CUDA(cudaGetDevice(&t->device));
    t->cap_log2 = cap_log2;
    t->chunk_chars = chunk_chars;
    t->epoch = 0;
    const size_t cap = (size_t)1 << cap_log2;
    cudaError_t e;
    if ((e = cudaMalloc(&t->tab, cap * sizeof(Entry))) != cudaSuccess ||
        (e = cudaMalloc(&t->dev, sizeof(TrieDev))) != cudaSuccess) {
        cudaFree(t->tab); cudaFree(t->dev);
        delete t;
        return cuda_fail(e, "cudaMalloc(trie)");
    }
    t->hs = (uint64_t *)t->tab;      // non-null: "store the chunk hash in the entry"
    int rc = b200r_trie_clear(t, nullptr);
    if (rc) { b200r_trie_destroy(t); return rc; }
    B200R_CUDA(cudaStreamSynchronize(nullptr));
    *out = t;
    return B200R_OK;
}

int b200r_trie_destroy(b200r_trie *t)
{
    if (!t) return B200R_OK;
    cudaFree(t->tab); cudaFree(t->dev);
    delete t;
    return B200R_OK;
}

int b200r_trie_clear(b200r_trie *t, void *stream)
{
    if (!t) { set_error("null trie"); return B200R_E_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    const size_t cap = (size_t)1 << t->cap_log2;
    B200R_CUDA(cudaMemsetAsync(t->tab, 0, cap * sizeof(Entry), st));
    B200R_CUDA(cudaMemsetAsync(t->dev, 0, sizeof(TrieDev), st));
    t->epoch = 0;
    return B200R_OK;
}

int b200r_trie_status(b200r_trie *t, void *stream, uint32_t *status_bits, int64_t *num_nodes,
                      uint64_t *root_mask, int64_t *capacity)
{
    if (!t) { set_error("null trie"); return B200R_E_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    if (num_nodes) {     // counted on demand: keeps a contended counter off the routing path
        const int64_t cap = (int64_t)1 << t->cap_log2;
        int blocks = (int)((cap + 255) / 256);
        if (blocks > 148 * 16) blocks = 148 * 16;
        B200R_CUDA(cudaMemsetAsync(&t->dev->num_nodes, 0, 8, st));
        k_count_nodes<<<blocks, 256, 0, st>>>(t->tab, cap, t->dev);
        B200R_LAUNCH_CHECK("k_count_nodes");
    }
    TrieDev h;
    B200R_CUDA(cudaMemcpyAsync(&h, t->dev, sizeof h, cudaMemcpyDeviceToHost, st));
    B200R_CUDA(cudaStreamSynchronize(st));
    if (status_bits) *status_bits = h.status;
    if (num_nodes) *num_nodes = (int64_t)h.num_nodes;
    if (root_mask) *root_mask = h.root_mask;
    if (capacity) *capacity = (int64_t)1 << t->cap_log2;
    return B200R_OK;
}

int b200r_trie_lookup_batch(b200r_trie *t, const uint64_t *d_hashes, const int32_t *d_chunk_base,
                            int64_t n, uint64_t avail_mask, int32_t *d_match_depth,
                            uint64_t *d_matched_mask, void *stream)
{
    if (!t || n < 0 || n > 0x7fffffff - 2048) { set_error("b200r_trie_lookup_batch: bad trie or n"); return B200R_E_INVALID; }
    if (n == 0) return B200R_OK;
    if (!d_hashes || !d_chunk_base || !d_match_depth || !d_matched_mask) { set_error("null pointer"); return B200R_E_INVALID; }
    const int blocks = (int)((n + WARPS_PER_CTA - 1) / WARPS_PER_CTA);
    { ProfScope ps__(PK_LOOKUP, (cudaStream_t)stream); g_launches += 1;
    k_lookup<<<blocks, WARPS_PER_CTA * 32, 0, (cudaStream_t)stream>>>(
        t->tab, t->cap_log2, d_hashes, d_chunk_base, (int)n, avail_mask, d_match_depth,
        d_matched_mask);
    B200R_LAUNCH_CHECK("k_lookup");
    }
    return B200R_OK;
}

int b200r_trie_insert_batch(b200r_trie *t, const uint64_t *d_hashes, const int32_t *d_chunk_base,
                            int64_t n, const int32_t *d_endpoint, void *stream)
{
    if (!t || n < 0 || n > 0x7fffffff - 2048) { set_error("b200r_trie_insert_batch: bad trie or n"); return B200R_E_INVALID; }
    if (n == 0) return B200R_OK;
    if (!d_hashes || !d_chunk_base || !d_endpoint) { set_error("null pointer"); return B200R_E_INVALID; }
    const int blocks = (int)((n + WARPS_PER_CTA - 1) / WARPS_PER_CTA);
    { ProfScope ps__(PK_INSERT, (cudaStream_t)stream); g_launches += 1;
    k_insert<<<blocks, WARPS_PER_CTA * 32, 0, (cudaStream_t)stream>>>(
        t->tab, t->hs, t->cap_log2, t->dev, d_hashes, d_chunk_base, (int)n, d_endpoint);
    B200R_LAUNCH_CHECK("k_insert");
    }
    return B200R_OK;
}

size_t b200r_prefix_route_scratch_bytes(int64_t n, int64_t max_chunks, uint32_t flags)
{
    return make_plan(n, max_chunks, flags, true).total;
}

// The routed batch is two stages that only meet in the scratch buffer:
//   hash stage   prep + chunk hashing: reads the prompt bytes, does NOT touch the table
//   route stage  claim / select / commit: reads the hashes, updates the table
// so a caller may run the hash stage of batch b+1 (HBM-bandwidth bound) on one stream
// while the route stage of batch b (latency bound) runs on another, with one scratch
// buffer per batch in flight.  Route stages must be issued in batch order.
int b200r_prefix_hash_stage(b200r_trie *t, const uint8_t *d_bytes, const int64_t *d_starts,
                            const int32_t *d_lens, const int32_t *d_nchars, int64_t n,
                            int64_t max_chunks, uint32_t flags, void *d_scratch,
                            size_t scratch_bytes, void *stream)
{
    if (!t) { set_error("null trie"); return B200R_E_INVALID; }
    int rc = check_prompt_args(d_bytes, d_starts, d_lens, n, max_chunks, t->chunk_chars);
    if (rc) return rc;
    if (n == 0) return B200R_OK;
    const Plan p = make_plan(n, max_chunks, flags, true);
    if (!d_scratch || scratch_bytes < p.total) {
        set_error("scratch too small: need %zu bytes, got %zu", p.total, scratch_bytes);
        return B200R_E_SCRATCH;
    }
    uint8_t *sc = (uint8_t *)d_scratch;
    return launch_hash(d_bytes, d_starts, d_lens, d_nchars, n, t->chunk_chars, flags,
                       (int32_t *)(sc + p.off_chunk_base), (uint64_t *)(sc + p.off_hashes),
                       max_chunks, sc, p, t->dev, (cudaStream_t)stream, false,
                       (int32_t *)(sc + p.off_chunk_req));
}

int b200r_prefix_route_stage(b200r_trie *t, int64_t n, int64_t max_chunks, uint32_t flags,
                             uint64_t avail_mask, uint64_t seed, uint64_t seq0,
                             int32_t *d_endpoint, int32_t *d_match_depth,
                             uint64_t *d_matched_mask, void *d_scratch, size_t scratch_bytes,
                             void *stream)
{
    if (!t || n < 0 || max_chunks < 0) { set_error("b200r_prefix_route_stage: bad argument"); return B200R_E_INVALID; }
    if (n == 0) return B200R_OK;
    if (!d_endpoint || !d_match_depth || !d_matched_mask) { set_error("null output pointer"); return B200R_E_INVALID; }
    const Plan p = make_plan(n, max_chunks, flags, true);
    if (!d_scratch || scratch_bytes < p.total) {
        set_error("scratch too small: need %zu bytes, got %zu", p.total, scratch_bytes);
        return B200R_E_SCRATCH;
    }
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t *sc = (uint8_t *)d_scratch;
    int32_t *chunk_base = (int32_t *)(sc + p.off_chunk_base);
    uint64_t *hashes = (uint64_t *)(sc + p.off_hashes);
    uint32_t *slots = (uint32_t *)(sc + p.off_slots);
    int32_t *leader = (int32_t *)(sc + p.off_leader);
    int32_t *choice = (int32_t *)(sc + p.off_choice);

    // batch epoch of the "first toucher" words: value = (epoch << 32) | ~request, merged
    // with atomicMax, so a newer batch overrides whatever older batches left and, inside
    // a batch, the smallest request index wins.  Zero (fresh table) means "never".
    if (t->epoch >= 0xFFFFFFF0u) {
        set_error("batch epoch exhausted; b200r_trie_clear or re-create the table");
        return B200R_E_INVALID;
    }
    t->epoch++;
    const unsigned long long tag = (unsigned long long)t->epoch << 32;

    const int blocks = (int)((n + WARPS_PER_CTA - 1) / WARPS_PER_CTA);
    const int thr = WARPS_PER_CTA * 32;
    const int tblocks = (int)((n + 255) / 256);
    { ProfScope ps__(PK_CLAIM, st); g_launches += 1;
    if (g_claim_by_tiles) {
        // one warp per 32-chunk tile (dense lanes), carry recomputed from the hashes
        FuseArgs fa = FuseArgs();
        fa.tab = t->tab; fa.hs = t->hs; fa.cap_log2 = t->cap_log2; fa.dev = t->dev;
        fa.avail = avail_mask; fa.tag = tag; fa.slots = slots;
        fa.chunk_req = (int32_t *)(sc + p.off_chunk_req);
        const int ntiles = (int)((max_chunks + HASH_WARPS * WTILE - 1) / (HASH_WARPS * WTILE));
        if (ntiles > 0)
            k_hash_tiles<2><<<ntiles, HASH_WARPS * 32, 0, st>>>(
                nullptr, nullptr, nullptr, nullptr, chunk_base, (const int32_t *)(sc + p.off_tile_first),
                nullptr, (int)n, t->chunk_chars, max_chunks, hashes, fa);
    } else {
        k_claim<<<blocks, thr, 0, st>>>(t->tab, t->hs, t->cap_log2, t->dev, hashes, chunk_base,
                                        (int)n, max_chunks, avail_mask, tag, slots, nullptr);
    }
    B200R_LAUNCH_CHECK("k_claim");
    }
    { ProfScope ps__(PK_SELECT, st); g_launches += 1;
    k_select<<<tblocks, 256, 0, st>>>(t->tab, chunk_base, (int)n, max_chunks, slots,
                                     avail_mask, tag, seed, seq0, leader, choice, d_match_depth,
                                     d_matched_mask, 0u, nullptr, nullptr, nullptr);
    B200R_LAUNCH_CHECK("k_select");
    }
    { ProfScope ps__(PK_COMMIT, st); g_launches += 1;
    k_commit<<<tblocks, 256, 0, st>>>(t->tab, t->dev, chunk_base, (int)n, max_chunks, slots, leader,
                                     choice, d_match_depth, d_endpoint, d_matched_mask, nullptr);
    B200R_LAUNCH_CHECK("k_commit");
    }
    return B200R_OK;
}

// Fused variant: prep stage (chunk counts, CSR, ragged tails; never touches the table)
// and fused stage (hash + path keys + claim in ONE kernel, then select and commit).
int b200r_prefix_prep_stage(b200r_trie *t, const uint8_t *d_bytes, const int64_t *d_starts,
                            const int32_t *d_lens, const int32_t *d_nchars, int64_t n,
                            int64_t max_chunks, uint32_t flags, void *d_scratch,
                            size_t scratch_bytes, void *stream)
{
    if (!t) { set_error("null trie"); return B200R_E_INVALID; }
    int rc = check_prompt_args(d_bytes, d_starts, d_lens, n, max_chunks, t->chunk_chars);
    if (rc) return rc;
    if (n == 0) return B200R_OK;
    const Plan p = make_plan(n, max_chunks, flags, true);
    if (!d_scratch || scratch_bytes < p.total) {
        set_error("scratch too small: need %zu bytes, got %zu", p.total, scratch_bytes);
        return B200R_E_SCRATCH;
    }
    uint8_t *sc = (uint8_t *)d_scratch;
    return launch_hash(d_bytes, d_starts, d_lens, d_nchars, n, t->chunk_chars, flags,
                       (int32_t *)(sc + p.off_chunk_base), (uint64_t *)(sc + p.off_hashes),
                       max_chunks, sc, p, t->dev, (cudaStream_t)stream, true);
}

int b200r_prefix_fused_stage(b200r_trie *t, const uint8_t *d_bytes, const int64_t *d_starts,
                             const int32_t *d_lens, const int32_t *d_nchars, int64_t n,
                             int64_t max_chunks, uint32_t flags, uint64_t avail_mask,
                             uint64_t seed, uint64_t seq0, int32_t *d_endpoint,
                             int32_t *d_match_depth, uint64_t *d_matched_mask, void *d_scratch,
                             size_t scratch_bytes, void *stream)
{
    if (!t) { set_error("null trie"); return B200R_E_INVALID; }
    int rc = check_prompt_args(d_bytes, d_starts, d_lens, n, max_chunks, t->chunk_chars);
    if (rc) return rc;
    if (n == 0) return B200R_OK;
    if (!d_endpoint || !d_match_depth || !d_matched_mask) { set_error("null output pointer"); return B200R_E_INVALID; }
    const Plan p = make_plan(n, max_chunks, flags, true);
    if (!d_scratch || scratch_bytes < p.total) {
        set_error("scratch too small: need %zu bytes, got %zu", p.total, scratch_bytes);
        return B200R_E_SCRATCH;
    }
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t *sc = (uint8_t *)d_scratch;
    int32_t *chunk_base = (int32_t *)(sc + p.off_chunk_base);
    uint64_t *hashes = (uint64_t *)(sc + p.off_hashes);
    uint32_t *slots = (uint32_t *)(sc + p.off_slots);
    int32_t *leader = (int32_t *)(sc + p.off_leader);
    int32_t *choice = (int32_t *)(sc + p.off_choice);
    const bool ascii = flags & B200R_F_ALL_ASCII;
    const int32_t *nchars = ascii ? nullptr : (d_nchars ? d_nchars : (const int32_t *)(sc + p.off_nchars));
    const int32_t *chunk_off = ascii ? nullptr : (const int32_t *)(sc + p.off_chunk_off);
    if (t->epoch >= 0xFFFFFFF0u) { set_error("batch epoch exhausted; b200r_trie_clear or re-create the table"); return B200R_E_INVALID; }
    t->epoch++;
    const unsigned long long tag = (unsigned long long)t->epoch << 32;
    FuseArgs fa;
    fa.tab = t->tab; fa.hs = t->hs; fa.cap_log2 = t->cap_log2; fa.dev = t->dev;
    fa.avail = avail_mask; fa.tag = tag; fa.slots = slots;
    fa.ticket = (unsigned int *)(sc + p.off_ticket);
    fa.lb_flag = (unsigned int *)(sc + p.off_lb_flag);
    fa.lb_part = (uint64_t *)(sc + p.off_lb_part);
    fa.lb_final = (uint64_t *)(sc + p.off_lb_final);
    const int ntiles = (int)((max_chunks + HASH_WARPS * WTILE - 1) / (HASH_WARPS * WTILE));
    if (ntiles > 0) {
        ProfScope ps__(PK_HASH, st); g_launches += 1;
        k_hash_tiles<1><<<ntiles, HASH_WARPS * 32, 0, st>>>(
            d_bytes, d_starts, d_lens, nchars, chunk_base, (const int32_t *)(sc + p.off_tile_first),
            chunk_off, (int)n, t->chunk_chars, max_chunks, hashes, fa);
        B200R_LAUNCH_CHECK("k_hash_tiles<fused>");
    }
    const int tblocks = (int)((n + 255) / 256);
    { ProfScope ps__(PK_SELECT, st); g_launches += 1;
    k_select<<<tblocks, 256, 0, st>>>(t->tab, chunk_base, (int)n, max_chunks, slots, avail_mask, tag,
                                     seed, seq0, leader, choice, d_match_depth, d_matched_mask, 0u,
                                     nullptr, nullptr, nullptr);
    B200R_LAUNCH_CHECK("k_select");
    }
    { ProfScope ps__(PK_COMMIT, st); g_launches += 1;
    k_commit<<<tblocks, 256, 0, st>>>(t->tab, t->dev, chunk_base, (int)n, max_chunks, slots, leader,
                                     choice, d_match_depth, d_endpoint, d_matched_mask, nullptr);
    B200R_LAUNCH_CHECK("k_commit");
    }
    return B200R_OK;
}

int b200r_prefix_route_batch_fused(b200r_trie *t, const uint8_t *d_bytes, const int64_t *d_starts,
                                   const int32_t *d_lens, const int32_t *d_nchars, int64_t n,
                                   int64_t max_chunks, uint32_t flags, uint64_t avail_mask,
                                   uint64_t seed, uint64_t seq0, int32_t *d_endpoint,
                                   int32_t *d_match_depth, uint64_t *d_matched_mask,
                                   void *d_scratch, size_t scratch_bytes, void *stream)
{
    int rc = b200r_prefix_prep_stage(t, d_bytes, d_starts, d_lens, d_nchars, n, max_chunks, flags,
                                     d_scratch, scratch_bytes, stream);
    if (rc) return rc;
    return b200r_prefix_fused_stage(t, d_bytes, d_starts, d_lens, d_nchars, n, max_chunks, flags,
                                    avail_mask, seed, seq0, d_endpoint, d_match_depth,
                                    d_matched_mask, d_scratch, scratch_bytes, stream);
}

int b200r_prefix_route_batch(b200r_trie *t, const uint8_t *d_bytes, const int64_t *d_starts,
                             const int32_t *d_lens, const int32_t *d_nchars, int64_t n,
                             int64_t max_chunks, uint32_t flags, uint64_t avail_mask,
                             uint64_t seed, uint64_t seq0, int32_t *d_endpoint,
                             int32_t *d_match_depth, uint64_t *d_matched_mask, void *d_scratch,
                             size_t scratch_bytes, void *stream)
{
    int rc = b200r_prefix_hash_stage(t, d_bytes, d_starts, d_lens, d_nchars, n, max_chunks, flags,
                                     d_scratch, scratch_bytes, stream);
    if (rc) return rc;
    return b200r_prefix_route_stage(t, n, max_chunks, flags, avail_mask, seed, seq0, d_endpoint,
                                    d_match_depth, d_matched_mask, d_scratch, scratch_bytes, stream);
}

// ---------------------------------------------------------------- multi-GPU
int b200r_shard_classify(b200r_trie *t, const uint8_t *d_bytes, const int64_t *d_starts,
                         const int32_t *d_lens, const int32_t *d_nchars, int64_t n,
                         int64_t max_chunks, uint32_t flags, uint64_t avail_mask,
                         void *d_records, uint32_t *d_rec_count, void *d_scratch,
                         size_t scratch_bytes, void *stream)
{
    if (!t) { set_error("null trie"); return B200R_E_INVALID; }
    int rc = check_prompt_args(d_bytes, d_starts, d_lens, n, max_chunks, t->chunk_chars);
    if (rc) return rc;
    if (!d_rec_count || (max_chunks > 0 && !d_records)) { set_error("null record buffer"); return B200R_E_INVALID; }
    const Plan p = make_plan(n, max_chunks, flags, true);
    if (!d_scratch || scratch_bytes < p.total) {
        set_error("scratch too small: need %zu bytes, got %zu", p.total, scratch_bytes);
        return B200R_E_SCRATCH;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (t->epoch >= 0xFFFFFFF0u) { set_error("batch epoch exhausted"); return B200R_E_INVALID; }
    t->epoch++;
    const uint32_t init[2] = { 0u, (uint32_t)n };       // {record count, shard size}
    B200R_CUDA(cudaMemcpyAsync(d_rec_count, init, 8, cudaMemcpyHostToDevice, st));
    if (n == 0) return B200R_OK;
    uint8_t *sc = (uint8_t *)d_scratch;
    int32_t *chunk_base = (int32_t *)(sc + p.off_chunk_base);
    uint64_t *hashes = (uint64_t *)(sc + p.off_hashes);
    uint32_t *slots = (uint32_t *)(sc + p.off_slots);
    rc = launch_hash(d_bytes, d_starts, d_lens, d_nchars, n, t->chunk_chars, flags, chunk_base,
                     hashes, max_chunks, sc, p, t->dev, st);
    if (rc) return rc;
    const int blocks = (int)((n + WARPS_PER_CTA - 1) / WARPS_PER_CTA);
    g_launches += 1;
    { ProfScope ps__(PK_CLASSIFY, st);
    k_shard_classify<<<blocks, WARPS_PER_CTA * 32, 0, st>>>(
        t->tab, t->cap_log2, hashes, chunk_base, (int)n, max_chunks, avail_mask, slots,
        (ShardRec *)d_records, d_rec_count);
    B200R_LAUNCH_CHECK("k_shard_classify");
    }
    return B200R_OK;
}

int b200r_shard_apply(b200r_trie *t, const void *d_records_all, const int32_t *d_meta,
                      int32_t world, int64_t stride, int32_t my_rank, int64_t n,
                      int64_t max_chunks, uint32_t flags, uint32_t *d_rec_slot, void *d_scratch,
                      void *stream)
{
    if (!t || world <= 0 || stride < 0 || my_rank < 0 || my_rank >= world) { set_error("b200r_shard_apply: bad argument"); return B200R_E_INVALID; }
    if (stride == 0) return B200R_OK;
    if (!d_records_all || !d_meta || !d_rec_slot || !d_scratch) { set_error("null pointer"); return B200R_E_INVALID; }
    const Plan p = make_plan(n, max_chunks, flags, true);
    uint8_t *sc = (uint8_t *)d_scratch;
    const int64_t work = (int64_t)world * stride;
    const unsigned long long tag = (unsigned long long)t->epoch << 32;
    g_launches += 1;
    k_shard_apply<<<(int)((work + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        t->tab, t->hs, t->cap_log2, t->dev, (const ShardRec *)d_records_all, d_meta, world, stride,
        my_rank, (const int32_t *)(sc + p.off_chunk_base), tag, (uint32_t *)(sc + p.off_slots),
        d_rec_slot);
    B200R_LAUNCH_CHECK("k_shard_apply");
    return B200R_OK;
}

int b200r_shard_select(b200r_trie *t, int64_t n, int64_t max_chunks, uint32_t flags,
                       uint64_t avail_mask, uint64_t seed, uint64_t seq0, uint32_t req_off,
                       int32_t *d_leader, int32_t *d_choice, int32_t *d_depth,
                       uint64_t *d_matched, void *d_scratch, void *stream)
{
    if (!t || n < 0) { set_error("b200r_shard_select: bad argument"); return B200R_E_INVALID; }
    if (n == 0) return B200R_OK;
    if (!d_leader || !d_choice || !d_depth || !d_matched || !d_scratch) { set_error("null pointer"); return B200R_E_INVALID; }
    const Plan p = make_plan(n, max_chunks, flags, true);
    uint8_t *sc = (uint8_t *)d_scratch;
    const unsigned long long tag = (unsigned long long)t->epoch << 32;
    g_launches += 1;
    k_select<<<(int)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        t->tab, (const int32_t *)(sc + p.off_chunk_base), (int)n, max_chunks,
        (const uint32_t *)(sc + p.off_slots), avail_mask, tag, seed, seq0, d_leader, d_choice,
        d_depth, d_matched, req_off, nullptr, nullptr, nullptr);
    B200R_LAUNCH_CHECK("k_select");
    return B200R_OK;
}

int b200r_shard_commit(b200r_trie *t, const void *d_records_all, const int32_t *d_meta,
                       int32_t world, int64_t stride, const uint32_t *d_rec_slot,
                       const int32_t *d_lcd_all, int64_t rstride, int32_t my_rank, int64_t n,
                       int32_t *d_endpoint_all, int32_t *d_endpoint, int32_t *d_match_depth,
                       uint64_t *d_matched, void *stream)
{
    if (!t || world <= 0 || my_rank < 0 || my_rank >= world || rstride < 0 || stride < 0) { set_error("b200r_shard_commit: bad argument"); return B200R_E_INVALID; }
    if (rstride == 0) return B200R_OK;
    if (!d_lcd_all || !d_meta || !d_endpoint_all) { set_error("null pointer"); return B200R_E_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t rwork = (int64_t)world * rstride;
    g_launches += 1;
    { ProfScope ps__(PK_RESOLVE, st);
    k_shard_resolve<<<(int)((rwork + 255) / 256), 256, 0, st>>>(t->dev, d_lcd_all, d_meta, world,
                                                              rstride, d_endpoint_all);
    B200R_LAUNCH_CHECK("k_shard_resolve");
    }
    const int64_t work = (int64_t)world * stride;
    if (work > 0) {
        g_launches += 1;
        { ProfScope ps__(PK_COMMIT, st);
        k_shard_commit<<<(int)((work + 255) / 256), 256, 0, st>>>(
            t->tab, (const ShardRec *)d_records_all, d_meta, world, stride, d_rec_slot,
            d_endpoint_all, d_lcd_all, rstride);
        B200R_LAUNCH_CHECK("k_shard_commit");
        }
    }
    if (n > 0) {
        if (!d_endpoint || !d_match_depth || !d_matched) { set_error("null output pointer"); return B200R_E_INVALID; }
        g_launches += 1;
        { ProfScope ps__(PK_RESOLVE, st);
        k_shard_finish<<<(int)((n + 255) / 256), 256, 0, st>>>(d_endpoint_all, d_lcd_all, (int)n, my_rank,
                                                             rstride, d_endpoint, d_match_depth,
                                                             d_matched);
        B200R_LAUNCH_CHECK("k_shard_finish");
        }
    }
    return B200R_OK;
}

// --------------------------------------------------- multi-GPU over peer memory
struct P2PLayout { size_t total, recs_off, lcd_off, rec_set, lcd_set; };

static P2PLayout p2p_layout(int world, int64_t cap, int64_t rcap)
{
    P2PLayout L;
    L.recs_off = P2P_FLAGS_BYTES;
    L.rec_set = (size_t)world * (size_t)cap * sizeof(ShardRec);
    L.lcd_off = L.recs_off + 2 * L.rec_set;
    L.lcd_set = (size_t)world * 3 * (size_t)rcap * 4;
    L.total = align_up(L.lcd_off + 2 * L.lcd_set);
    return L;
}

static int p2p_check(int world, int rank, int64_t cap, int64_t rcap, const void *bases)
{
    if (world <= 0 || world > P2P_MAX_WORLD || rank < 0 || rank >= world || cap <= 0 || rcap <= 0 ||
        (rcap & 3) || !bases) {
        set_error("p2p: need 1 <= world <= %d, 0 <= rank < world, cap > 0, rcap > 0 and a multiple "
                  "of 4, peer base table", P2P_MAX_WORLD);
        return B200R_E_INVALID;
    }
    return B200R_OK;
}

int b200r_p2p_layout(int32_t world, int64_t cap, int64_t rcap, size_t out[3])
{
    if (world <= 0 || world > P2P_MAX_WORLD || cap <= 0 || rcap <= 0 || !out) { set_error("b200r_p2p_layout: bad argument"); return B200R_E_INVALID; }
    const P2PLayout L = p2p_layout(world, cap, rcap);
    out[0] = L.total; out[1] = L.recs_off; out[2] = L.lcd_off;
    return B200R_OK;
}

int b200r_p2p_classify(b200r_trie *t, const uint8_t *d_bytes, const int64_t *d_starts,
                       const int32_t *d_lens, const int32_t *d_nchars, int64_t n,
                       int64_t max_chunks, uint32_t flags, uint64_t avail_mask,
                       const uint64_t *d_peer_bases, int32_t world, int32_t rank, int64_t cap,
                       int64_t rcap, uint32_t epoch, void *d_recs_local, uint32_t *d_rec_count,
                       void *d_scratch, size_t scratch_bytes, void *stream)
{
    int rc = p2p_check(world, rank, cap, rcap, d_peer_bases);
    if (rc) return rc;
    if (max_chunks > cap || n > rcap) { set_error("p2p: batch (n=%lld, max_chunks=%lld) exceeds the exchange capacity (rcap=%lld, cap=%lld)", (long long)n, (long long)max_chunks, (long long)rcap, (long long)cap); return B200R_E_INVALID; }
    rc = b200r_shard_classify(t, d_bytes, d_starts, d_lens, d_nchars, n, max_chunks, flags,
                              avail_mask, d_recs_local, d_rec_count, d_scratch, scratch_bytes, stream);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const P2PLayout L = p2p_layout(world, cap, rcap);
    const int set = (int)(epoch & 1);
    const size_t dst = L.recs_off + (size_t)set * L.rec_set + (size_t)rank * (size_t)cap * sizeof(ShardRec);
    g_launches += 2;
    { ProfScope ps__(PK_PUSH, st);
    k_p2p_push<<<148 * 4, 256, 0, st>>>((const uint4 *)d_recs_local, d_rec_count, 0, 2,
                                        (const unsigned long long *)d_peer_bases, world, dst);
    B200R_LAUNCH_CHECK("k_p2p_push");
    }
    { ProfScope ps__(PK_PUSH, st);
    k_p2p_publish<<<1, P2P_MAX_WORLD, 0, st>>>((const unsigned long long *)d_peer_bases, world, rank,
                                               set, epoch, 0, d_rec_count, (unsigned)n);
    B200R_LAUNCH_CHECK("k_p2p_publish");
    }
    return B200R_OK;
}

int b200r_p2p_route(b200r_trie *t, int64_t n, int64_t max_chunks, uint32_t flags,
                    uint64_t avail_mask, uint64_t seed, const uint64_t *d_peer_bases,
                    void *d_my_base, int32_t world, int32_t rank, int64_t cap, int64_t rcap,
                    uint32_t epoch, void *d_dyn, uint32_t *d_rec_slot, int32_t *d_lcd_local,
                    uint64_t *d_matched, void *d_scratch, void *stream)
{
    int rc = p2p_check(world, rank, cap, rcap, d_peer_bases);
    if (rc) return rc;
    if (!t || !d_my_base || !d_dyn || !d_rec_slot || !d_lcd_local || !d_scratch || (n > 0 && !d_matched)) { set_error("b200r_p2p_route: null pointer"); return B200R_E_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    const P2PLayout L = p2p_layout(world, cap, rcap);
    const int set = (int)(epoch & 1);
    uint8_t *base = (uint8_t *)d_my_base;
    P2PDyn *dyn = (P2PDyn *)d_dyn;
    const ShardRec *recs_all = (const ShardRec *)(base + L.recs_off + (size_t)set * L.rec_set);
    const Plan p = make_plan(n, max_chunks, flags, true);
    uint8_t *sc = (uint8_t *)d_scratch;
    const unsigned long long tag = (unsigned long long)t->epoch << 32;
    g_launches += 1;
    { ProfScope ps__(PK_WAIT, st);
    k_p2p_wait_meta<<<1, P2P_MAX_WORLD, 0, st>>>((const P2PFlags *)base, set, epoch, world, rank, dyn, t->dev);
    B200R_LAUNCH_CHECK("k_p2p_wait_meta");
    }
    const int64_t work = (int64_t)world * cap;
    g_launches += 1;
    { ProfScope ps__(PK_APPLY, st);
    k_shard_apply<<<(int)((work + 255) / 256), 256, 0, st>>>(
        t->tab, t->hs, t->cap_log2, t->dev, recs_all, dyn->meta, world, cap, rank,
        (const int32_t *)(sc + p.off_chunk_base), tag, (uint32_t *)(sc + p.off_slots), d_rec_slot);
    B200R_LAUNCH_CHECK("k_shard_apply");
    }
    if (n > 0) {
        g_launches += 1;
        { ProfScope ps__(PK_SELECT, st);
        k_select<<<(int)((n + 255) / 256), 256, 0, st>>>(
            t->tab, (const int32_t *)(sc + p.off_chunk_base), (int)n, max_chunks,
            (const uint32_t *)(sc + p.off_slots), avail_mask, tag, seed, 0, d_lcd_local,
            d_lcd_local + rcap, d_lcd_local + 2 * rcap, d_matched, 0u, dyn, nullptr, nullptr);
        B200R_LAUNCH_CHECK("k_select");
        }
    }
    const size_t dst = L.lcd_off + (size_t)set * L.lcd_set + (size_t)rank * 3 * (size_t)rcap * 4;
    g_launches += 2;
    { ProfScope ps__(PK_PUSH, st);
    k_p2p_push<<<148, 256, 0, st>>>((const uint4 *)d_lcd_local, nullptr, 3 * rcap / 4, 1,
                                    (const unsigned long long *)d_peer_bases, world, dst);
    B200R_LAUNCH_CHECK("k_p2p_push(lcd)");
    }
    { ProfScope ps__(PK_PUSH, st);
    k_p2p_publish<<<1, P2P_MAX_WORLD, 0, st>>>((const unsigned long long *)d_peer_bases, world, rank,
                                               set, epoch, 1, nullptr, (unsigned)n);
    B200R_LAUNCH_CHECK("k_p2p_publish(lcd)");
    }
    return B200R_OK;
}

int b200r_p2p_commit(b200r_trie *t, int64_t n, void *d_my_base, int32_t world, int32_t rank,
                     int64_t cap, int64_t rcap, uint32_t epoch, void *d_dyn,
                     const uint32_t *d_rec_slot, int32_t *d_endpoint_all, int32_t *d_endpoint,
                     int32_t *d_match_depth, uint64_t *d_matched, void *stream)
{
    int rc = p2p_check(world, rank, cap, rcap, d_my_base);
    if (rc) return rc;
    if (!t || !d_dyn || !d_rec_slot || !d_endpoint_all) { set_error("b200r_p2p_commit: null pointer"); return B200R_E_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    const P2PLayout L = p2p_layout(world, cap, rcap);
    const int set = (int)(epoch & 1);
    uint8_t *base = (uint8_t *)d_my_base;
    P2PDyn *dyn = (P2PDyn *)d_dyn;
    const int32_t *lcd_all = (const int32_t *)(base + L.lcd_off + (size_t)set * L.lcd_set);
    g_launches += 1;
    { ProfScope ps__(PK_WAIT, st);
    k_p2p_wait_lcd<<<1, P2P_MAX_WORLD, 0, st>>>((const P2PFlags *)base, set, epoch, world, t->dev);
    B200R_LAUNCH_CHECK("k_p2p_wait_lcd");
    }
    rc = b200r_shard_commit(t, base + L.recs_off + (size_t)set * L.rec_set, dyn->meta, world, cap,
                            d_rec_slot, lcd_all, rcap, rank, n, d_endpoint_all, d_endpoint,
                            d_match_depth, d_matched, stream);
    if (rc) return rc;
    g_launches += 1;
    k_p2p_advance<<<1, 1, 0, st>>>(dyn);
    B200R_LAUNCH_CHECK("k_p2p_advance");
    return B200R_OK;
}

// ------------------------------------------------------ multi-GPU, partitioned
struct PartLayout {
    size_t symm_total, hash_off, info_off, res_off, hash_set, info_set, res_set;
    // local work buffer
    size_t work_total, w_dyn, w_blk_cnt, w_blk_off, w_totals, w_dest_slot, w_req_dst, w_hash_stage, w_info_stage, w_c_owned, w_pos_owned,
        w_ep, w_depth, w_matched, w_plan;
    int64_t n_max, c_max;
    int nblk_max;
};

static PartLayout part_layout(int world, int64_t cap, int64_t rcap)
{
    PartLayout L;
    L.hash_off = 4096;
    L.hash_set = (size_t)world * cap * 8;
    L.info_off = L.hash_off + 2 * L.hash_set;
    L.info_set = (size_t)world * rcap * 8;
    L.res_off = L.info_off + 2 * L.info_set;
    L.res_set = (size_t)world * rcap * sizeof(PartRes);
    L.symm_total = align_up(L.res_off + 2 * L.res_set);
    L.n_max = (int64_t)world * rcap;
    L.c_max = (int64_t)world * cap;
    L.nblk_max = (int)((rcap + PART_BLOCK - 1) / PART_BLOCK);
    size_t o = 0;
    L.w_dyn = o;       o += align_up(sizeof(PartDyn));
    L.w_blk_cnt = o;   o += align_up((size_t)L.nblk_max * 2 * world * 4);
    L.w_blk_off = o;   o += align_up((size_t)L.nblk_max * 2 * world * 4);
    L.w_totals = o;    o += align_up((size_t)4 * world * 4);
    L.w_dest_slot = o; o += 2 * align_up((size_t)rcap * 4);     // one per batch parity
    L.w_req_dst = o;   o += align_up((size_t)rcap * 4);
    L.w_hash_stage = o; o += align_up((size_t)(cap + 2 * world + 2) * 8);
    L.w_info_stage = o; o += align_up((size_t)(rcap + 2 * world + 2) * 8);
    L.w_c_owned = o;   o += align_up((size_t)L.n_max * 4);
    L.w_pos_owned = o; o += align_up((size_t)L.n_max * 4);
    L.w_ep = o;        o += align_up((size_t)L.n_max * 4);
    L.w_depth = o;     o += align_up((size_t)L.n_max * 4);
    L.w_matched = o;   o += align_up((size_t)L.n_max * 8);
    L.w_plan = o;      o += make_plan(L.n_max, L.c_max, B200R_F_ALL_ASCII, true).total;
    L.work_total = o + 256;
    return L;
}

static int part_check(int world, int rank, int64_t cap, int64_t rcap, const void *a, const void *b)
{
    if (world <= 0 || world > P2P_MAX_WORLD || rank < 0 || rank >= world || cap <= 0 || rcap <= 0 ||
        (cap & 1) || (rcap & 1) ||
        rcap >= (1 << 24) || (int64_t)world * cap > 0x7fffffff - 4096 || !a || !b) {
        set_error("partitioned routing: need 1 <= world <= %d, 0 <= rank < world, 0 < rcap < 2^24, "
                  "cap and rcap even, world * cap < 2^31, non-null buffers", P2P_MAX_WORLD);
        return B200R_E_INVALID;
    }
    return B200R_OK;
}

int b200r_part_layout(int32_t world, int64_t cap, int64_t rcap, size_t out[2])
{
    if (world <= 0 || world > P2P_MAX_WORLD || cap <= 0 || rcap <= 0 || !out) { set_error("b200r_part_layout: bad argument"); return B200R_E_INVALID; }
    const PartLayout L = part_layout(world, cap, rcap);
    out[0] = L.symm_total; out[1] = L.work_total;
    return B200R_OK;
}

int b200r_part_send(b200r_trie *t, const uint8_t *d_bytes, const int64_t *d_starts,
                    const int32_t *d_lens, const int32_t *d_nchars, int64_t n, int64_t max_chunks,
                    uint32_t flags, const uint64_t *d_peer_bases, int32_t world, int32_t rank,
                    int64_t cap, int64_t rcap, uint32_t epoch, void *d_work, void *d_scratch,
                    size_t scratch_bytes, void *stream)
{
    int rc = part_check(world, rank, cap, rcap, d_peer_bases, d_work);
    if (rc) return rc;
    if (max_chunks > cap || n > rcap) { set_error("partitioned routing: batch (n=%lld, max_chunks=%lld) exceeds the exchange capacity (rcap=%lld, cap=%lld)", (long long)n, (long long)max_chunks, (long long)rcap, (long long)cap); return B200R_E_INVALID; }
    rc = b200r_prefix_hash_stage(t, d_bytes, d_starts, d_lens, d_nchars, n, max_chunks, flags,
                                 d_scratch, scratch_bytes, stream);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const PartLayout L = part_layout(world, cap, rcap);
    const Plan p = make_plan(n, max_chunks, flags, true);
    uint8_t *sc = (uint8_t *)d_scratch, *wk = (uint8_t *)d_work;
    const int32_t *chunk_base = (const int32_t *)(sc + p.off_chunk_base);
    const uint64_t *hashes = (const uint64_t *)(sc + p.off_hashes);
    int32_t *blk_cnt = (int32_t *)(wk + L.w_blk_cnt), *blk_off = (int32_t *)(wk + L.w_blk_off);
    int32_t *totals = (int32_t *)(wk + L.w_totals);
    const int set = (int)(epoch & 1);
    const int nblk = (int)((n + PART_BLOCK - 1) / PART_BLOCK);
    ProfScope ps__(PK_PUSH, st);
    if (nblk > 0) {
        g_launches += 3;
        k_part_count<<<nblk, PART_BLOCK, 0, st>>>(chunk_base, hashes, (int)n, world, blk_cnt);
        B200R_LAUNCH_CHECK("k_part_count");
        k_part_scan<<<1, 1024, 0, st>>>(blk_cnt, nblk, world, blk_off, totals);
        B200R_LAUNCH_CHECK("k_part_scan");
        uint32_t *dest_slot = (uint32_t *)(wk + L.w_dest_slot + (size_t)set * align_up((size_t)rcap * 4));
        uint32_t *req_dst = (uint32_t *)(wk + L.w_req_dst);
        uint64_t *hash_stage = (uint64_t *)(wk + L.w_hash_stage);
        unsigned long long *info_stage = (unsigned long long *)(wk + L.w_info_stage);
        k_part_push<<<nblk, PART_BLOCK, 0, st>>>(chunk_base, hashes, (int)n, world, blk_off, totals,
                                                 info_stage, dest_slot, req_dst);
        B200R_LAUNCH_CHECK("k_part_push");
        g_launches += 2;
        k_part_copy<<<(unsigned)std::min<int64_t>((n + WARPS_PER_CTA - 1) / WARPS_PER_CTA, 148 * 8),
                      WARPS_PER_CTA * 32, 0, st>>>(chunk_base, hashes, (int)n, req_dst, hash_stage);
        B200R_LAUNCH_CHECK("k_part_copy");
        k_part_bulk<<<148 * 2, 256, 0, st>>>(hash_stage, info_stage, totals, world, rank,
                                             (const unsigned long long *)d_peer_bases,
                                             L.hash_off + (size_t)set * L.hash_set,
                                             L.info_off + (size_t)set * L.info_set, cap, rcap);
        B200R_LAUNCH_CHECK("k_part_bulk");
    } else {
        B200R_CUDA(cudaMemsetAsync(totals, 0, (size_t)4 * world * 4, st));
    }
    g_launches += 1;
    k_part_publish_in<<<1, P2P_MAX_WORLD, 0, st>>>((const unsigned long long *)d_peer_bases, world, rank,
                                                   set, epoch, totals, (unsigned)n);
    B200R_LAUNCH_CHECK("k_part_publish_in");
    return B200R_OK;
}

int b200r_part_route(b200r_trie *t, uint64_t avail_mask, uint64_t seed,
                     const uint64_t *d_peer_bases, void *d_my_base, int32_t world, int32_t rank,
                     int64_t cap, int64_t rcap, uint32_t epoch, void *d_work, void *stream)
{
    int rc = part_check(world, rank, cap, rcap, d_peer_bases, d_work);
    if (rc) return rc;
    if (!t || !d_my_base) { set_error("b200r_part_route: null pointer"); return B200R_E_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    const PartLayout L = part_layout(world, cap, rcap);
    const Plan p = make_plan(L.n_max, L.c_max, B200R_F_ALL_ASCII, true);
    const int set = (int)(epoch & 1);
    uint8_t *base = (uint8_t *)d_my_base, *wk = (uint8_t *)d_work, *sc = wk + L.w_plan;
    PartDyn *dyn = (PartDyn *)(wk + L.w_dyn);
    int32_t *c_owned = (int32_t *)(wk + L.w_c_owned);
    uint32_t *pos_owned = (uint32_t *)(wk + L.w_pos_owned);
    int32_t *ep = (int32_t *)(wk + L.w_ep), *depth = (int32_t *)(wk + L.w_depth);
    uint64_t *matched = (uint64_t *)(wk + L.w_matched);
    int32_t *chunk_base = (int32_t *)(sc + p.off_chunk_base);
    uint64_t *hashes = (uint64_t *)(sc + p.off_hashes);
    uint32_t *slots = (uint32_t *)(sc + p.off_slots);
    int32_t *leader = (int32_t *)(sc + p.off_leader), *choice = (int32_t *)(sc + p.off_choice);
    int32_t *tile_first = (int32_t *)(sc + p.off_tile_first), *block_sums = (int32_t *)(sc + p.off_block_sums);

    if (t->epoch >= 0xFFFFFFF0u) { set_error("batch epoch exhausted"); return B200R_E_INVALID; }
    t->epoch++;
    const unsigned long long tag = (unsigned long long)t->epoch << 32;
    const int32_t *n_dev = &dyn->n_owned;

    { ProfScope ps__(PK_WAIT, st); g_launches += 1;
    k_part_wait_meta<<<1, P2P_MAX_WORLD, 0, st>>>((const PartFlags *)base, set, epoch, world, dyn, t->dev);
    B200R_LAUNCH_CHECK("k_part_wait_meta");
    }
    { ProfScope ps__(PK_PREP, st); g_launches += 3;
    k_part_gather<<<148 * 4, 256, 0, st>>>(base, L.hash_off + (size_t)set * L.hash_set,
                                           L.info_off + (size_t)set * L.info_set, cap, rcap, world, dyn,
                                           c_owned, pos_owned, hashes);
    B200R_LAUNCH_CHECK("k_part_gather");
    const int nblk = (int)((L.n_max + PREP_BLOCK - 1) / PREP_BLOCK);
    k_prep_sums<<<nblk, PREP_BLOCK, 0, st>>>(c_owned, nullptr, (int)L.n_max, 1, block_sums, n_dev, nullptr);
    B200R_LAUNCH_CHECK("k_prep_sums(owned)");
    k_prep_scan<<<nblk, PREP_BLOCK, 0, st>>>(c_owned, nullptr, (int)L.n_max, 1, block_sums, chunk_base,
                                             tile_first, L.c_max, t->dev, nullptr, nullptr, nullptr, n_dev, nullptr);
    B200R_LAUNCH_CHECK("k_prep_scan(owned)");
    }
    // grids are estimates (about two shards' worth); the kernels stride over the rest
    const int64_t est = 2 * rcap < L.n_max ? 2 * rcap : L.n_max;
    const int wblocks = (int)((est + WARPS_PER_CTA - 1) / WARPS_PER_CTA);
    const int tblocks = (int)((est + 255) / 256);
    { ProfScope ps__(PK_CLAIM, st); g_launches += 1;
    k_claim<<<wblocks, WARPS_PER_CTA * 32, 0, st>>>(t->tab, t->hs, t->cap_log2, t->dev, hashes,
                                                    chunk_base, 0, L.c_max, avail_mask, tag, slots, n_dev);
    B200R_LAUNCH_CHECK("k_claim(owned)");
    }
    { ProfScope ps__(PK_SELECT, st); g_launches += 1;
    k_select<<<tblocks, 256, 0, st>>>(t->tab, chunk_base, 0, L.c_max, slots, avail_mask, tag, seed, 0,
                                     leader, choice, depth, matched, 0u, &dyn->base, pos_owned, n_dev);
    B200R_LAUNCH_CHECK("k_select(owned)");
    }
    { ProfScope ps__(PK_COMMIT, st); g_launches += 1;
    k_commit<<<tblocks, 256, 0, st>>>(t->tab, t->dev, chunk_base, 0, L.c_max, slots, leader, choice,
                                     depth, ep, matched, n_dev);
    B200R_LAUNCH_CHECK("k_commit(owned)");
    }
    { ProfScope ps__(PK_PUSH, st); g_launches += 2;
    k_part_push_results<<<148 * 2, 256, 0, st>>>(dyn, world, rank, ep, depth, matched,
                                                 (const unsigned long long *)d_peer_bases,
                                                 L.res_off + (size_t)set * L.res_set, rcap);
    B200R_LAUNCH_CHECK("k_part_push_results");
    k_part_publish_res<<<1, P2P_MAX_WORLD, 0, st>>>((const unsigned long long *)d_peer_bases, world,
                                                    rank, set, epoch, dyn);
    B200R_LAUNCH_CHECK("k_part_publish_res");
    }
    return B200R_OK;
}

int b200r_part_finish(b200r_trie *t, int64_t n, void *d_my_base, int32_t world, int32_t rank,
                      int64_t cap, int64_t rcap, uint32_t epoch, void *d_work, int32_t *d_endpoint,
                      int32_t *d_match_depth, uint64_t *d_matched, void *stream)
{
    int rc = part_check(world, rank, cap, rcap, d_my_base, d_work);
    if (rc) return rc;
    if (!t || (n > 0 && (!d_endpoint || !d_match_depth || !d_matched))) { set_error("b200r_part_finish: null pointer"); return B200R_E_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    const PartLayout L = part_layout(world, cap, rcap);
    const int set = (int)(epoch & 1);
    uint8_t *base = (uint8_t *)d_my_base, *wk = (uint8_t *)d_work;
    { ProfScope ps__(PK_WAIT, st); g_launches += 1;
    k_part_wait_res<<<1, P2P_MAX_WORLD, 0, st>>>((const PartFlags *)base, set, epoch, world, t->dev,
                                                 (PartDyn *)(wk + L.w_dyn));
    B200R_LAUNCH_CHECK("k_part_wait_res");
    }
    if (n > 0) {
        ProfScope ps__(PK_RESOLVE, st); g_launches += 1;
        k_part_scatter<<<(int)((n + 255) / 256), 256, 0, st>>>(
            (const uint32_t *)(wk + L.w_dest_slot + (size_t)set * align_up((size_t)rcap * 4)), base,
            L.res_off + (size_t)set * L.res_set, rcap,
            (int)n, d_endpoint, d_match_depth, d_matched);
        B200R_LAUNCH_CHECK("k_part_scatter");
    }
    return B200R_OK;
}

int b200r_trie_dump(b200r_trie *t, uint64_t *d_out, int64_t max_nodes, int64_t *count,
                    uint64_t *root_mask, void *stream)
{
    if (!t || max_nodes < 0 || (max_nodes > 0 && !d_out)) { set_error("b200r_trie_dump: bad argument"); return B200R_E_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    B200R_CUDA(cudaMemsetAsync(&t->dev->dump_count, 0, 8, st));
    const int64_t cap = (int64_t)1 << t->cap_log2;
    int blocks = (int)((cap + 255) / 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_dump<<<blocks, 256, 0, st>>>(t->tab, t->hs, cap, t->dev, d_out, max_nodes);
    B200R_LAUNCH_CHECK("k_dump");
    TrieDev h;
    B200R_CUDA(cudaMemcpyAsync(&h, t->dev, sizeof h, cudaMemcpyDeviceToHost, st));
    B200R_CUDA(cudaStreamSynchronize(st));
    if (count) *count = (int64_t)h.dump_count;
    if (root_mask) *root_mask = h.root_mask;
    return B200R_OK;
}

int b200r_trie_apply_remote(b200r_trie *t, const uint64_t *d_entries, int64_t count,
                            uint64_t root_mask_or, void *stream)
{
    if (!t || count < 0 || (count > 0 && !d_entries)) { set_error("b200r_trie_apply_remote: bad argument"); return B200R_E_INVALID; }
    const int64_t work = count > 0 ? count : 1;
    const int blocks = (int)((work + 255) / 256);
    k_apply_remote<<<blocks, 256, 0, (cudaStream_t)stream>>>(t->tab, t->hs, t->cap_log2, t->dev, d_entries,
                                                           count, root_mask_or);
    B200R_LAUNCH_CHECK("k_apply_remote");
    return B200R_OK;
}

}  // extern "C"
