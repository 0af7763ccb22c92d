// prefix.cu -- batched prefix-aware routing on sm_100a.
//
// Replaces, behind the C ABI of include/b200_router.h, the reference's
//   HashTrie._chunk_and_hash / longest_prefix_match / insert
//       (src/vllm_router/prefix/hashtrie.py:45-103)
//   PrefixAwareRouter.route_request lookup -> choice -> insert
//       (src/vllm_router/routers/routing_logic.py:414-423)
//
// Kernel pipeline of one routed batch (no host sync):
//   hash stage   k_prep_sums / k_prep_scan   chunk counts -> CSR chunk_base, tile map
//                [k_count_chars, k_utf8_bounds   only for batches that are not pure ASCII]
//                k_hash_chunks               XXH64 of every chunk, one thread per chunk,
//                                            bytes staged by cp.async into swizzled smem
//   route stage  claim     k_claim_thread (a thread per request; short prompts) or
//                          k_walk<W_CLAIM> (8/16/32 lanes per request; long prompts), trie.cuh:
//                          stored runs of the node log are followed, missing paths appended as
//                          one run, "first toucher" atomicMax on nodes not usable before
//                k_select  per request: deepest usable node (bisection), root draw or leader
//                k_commit  resolve leader chains, OR the endpoint into the nodes below the
//                          match depth (k_commit_warp for long prompts)
// Around them: the three-stream batch pipeline (b200r_pipe_*), the one-request latency path
// (k_route_one), status / dump / restore / sweeps, and the C ABI.  The partitioned multi-GPU
// path is multigpu.cu; its send side hooks into k_prep_scan (PartPlace) and k_hash_chunks<true>.
#include <cstdarg>
#include <mutex>
#include <algorithm>
#include <vector>

#include "internal.cuh"
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

// ------------------------------------------------------------- profiling (internal.cuh)
int g_claim_mode = 0;          // b200r_set_claim_mode: 0 auto, else claim_kind_for's value
int g_part_staged = 0;
int g_fuse_sc = getenv("B200R_NO_FUSED_COMMIT") ? 0 : 1;   // select + commit in one launch where possible
int g_hash_mode = 1;           // b200r_set_hash_mode: 0 = persistent streaming kernel, 1 = one tile per warp
bool g_prof_on = false;
std::vector<ProfSpan> g_spans;
long long g_launches = 0;

constexpr int PREP_BLOCK = 256;    // requests per CTA of the prep kernels

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
            int32_t *__restrict__ block_sums, const int32_t *__restrict__ n_dev)
{
    __shared__ int s_red[32];
    if (n_dev) {                    // batch size known only on the device (partitioned multi-GPU):
        n = *n_dev;                 // the grid covers the capacity, blocks past the batch leave
        if ((int)blockIdx.x * PREP_BLOCK >= n) return;
    }
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
            uint64_t *__restrict__ hashes, const int32_t *__restrict__ n_dev, const PartPlace pp)
{
    __shared__ int s_red[32];
    __shared__ int s_wsum[32];
    if (n_dev) {
        n = *n_dev;
        if (n == 0 && blockIdx.x == 0 && threadIdx.x == 0) chunk_base[0] = 0;
        if ((int)blockIdx.x * PREP_BLOCK >= n) return;
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
    bool have_tail = false;                         // ragged last chunk hashed here
    uint64_t tailh = 0;
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
            if (tile_first)
                for (int t = t0; (int64_t)t * 32 < lim; t++) tile_first[t] = i;
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
                    have_tail = true;
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
                        tailh = b200r_th::xxh64_tail_words(wv, tail);
                    } else {
                        tailh = xxh64_global(tp, tail);
                    }
                    if (!pp.world) hashes[gl] = tailh;
                }
            }
        }
    }
    if (!pp.world) return;
    // ---- partitioned multi-GPU: place the request in its owner's inbound buffer (PartPlace)
    const int lane_ = tid & 31;
    int o = -1;
    if (i < n) {
        if (c == 0) o = 0;                          // empty prompt: touches only a root
        else {
            uint64_t h0;
            const int ln = lens[i];
            if (c == 1 && have_tail) h0 = tailh;
            else {
                const uint8_t *q = bytes + starts[i];
                const int l0 = ln < cc ? ln : cc;
                if (l0 == 128 && (((uintptr_t)q) & 15) == 0) {
                    uint64_t v1 = XP1 + XP2, v2 = XP2, v3 = 0, v4 = 0 - XP1;
                    uint4 qa[8];
#pragma unroll
                    for (int u = 0; u < 8; u++) qa[u] = __ldg((const uint4 *)q + u);
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        v1 = xround(v1, u64_of(qa[2 * u].x, qa[2 * u].y));
                        v2 = xround(v2, u64_of(qa[2 * u].z, qa[2 * u].w));
                        v3 = xround(v3, u64_of(qa[2 * u + 1].x, qa[2 * u + 1].y));
                        v4 = xround(v4, u64_of(qa[2 * u + 1].z, qa[2 * u + 1].w));
                    }
                    uint64_t h = rotl<1>(v1) + rotl<7>(v2) + rotl<12>(v3) + rotl<18>(v4);
                    h = xmerge(h, v1); h = xmerge(h, v2); h = xmerge(h, v3); h = xmerge(h, v4);
                    h0 = xavalanche(h + 128);
                } else {
                    h0 = xxh64_global(q, l0);
                }
            }
            o = part_owner_of(h0, pp.world);
        }
    }
    uint32_t rreq = 0, rchunk = 0;
    for (int d = 0; d < pp.world; d++) {
        const unsigned m = __ballot_sync(0xffffffffu, o == d);
        if (!m) continue;                                   // warp-uniform
        const int v = (o == d) ? c : 0;
        int inc2 = v;
        for (int sft = 1; sft < 32; sft <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc2, sft);
            if (lane_ >= sft) inc2 += t;
        }
        const int tot = __shfl_sync(0xffffffffu, inc2, 31);
        unsigned long long bs = 0;
        if (lane_ == 0)
            bs = atomicAdd(&pp.totals[d], ((unsigned long long)__popc(m) << 32) | (unsigned long long)(uint32_t)tot);
        bs = __shfl_sync(0xffffffffu, bs, 0);
        if (o == d) {
            rreq = (uint32_t)(bs >> 32) + (uint32_t)__popc(m & ((1u << lane_) - 1));
            rchunk = (uint32_t)bs + (uint32_t)(inc2 - v);
        }
    }
    if (o >= 0) {
        pp.dest_slot[i] = ((uint32_t)o << 24) | rreq;
        const unsigned long long pb = pp.peer_bases[o];
        ((unsigned long long *)(uintptr_t)(pb + pp.info_off))[(size_t)pp.my_rank * pp.rcap + rreq] =
            (unsigned long long)(uint32_t)i | ((unsigned long long)rchunk << 32);
        const unsigned long long op = pb + pp.hash_off + ((size_t)pp.my_rank * pp.cap + rchunk) * 8;
        pp.out_ptr[i] = op;
        if (have_tail) ((uint64_t *)(uintptr_t)op)[c - 1] = tailh;
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

// ---------------------------------------------------------------------------
// k_hash_chunks: one thread per chunk, and every WARP is an independent worker
// on its own tile of 32 consecutive chunks -- no block-wide barrier anywhere, so
// the 40 resident warps of an SM (5 CTAs x 8 warps, 4 KB of staging each) are
// 40 independent latency chains and ~160 KB of prompt bytes stay in flight.
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
// CTAs per SM the hashing kernel is compiled for.  Six fit (40 registers), and alone the kernel
// is as fast with six as with five; but six fill the register file, and the route stage's
// kernels of the previous batch -- short, latency-bound, on a higher-priority stream -- then
// wait for hashing CTAs to retire before they can start.  With five (48 registers) they find
// room at once: the pipelined step went 74 -> 67 us with the hashing kernel itself unchanged.
constexpr int HASH_CTAS_PER_SM = 5;

struct ReqRec { int b, b1, ln, nc; long long st; int r; };

// ASCII: the batch is declared pure ASCII (code points = bytes): no nchars, no chunk_off, and a
// request's last chunk is recognised by its byte range alone.  Chunk indices are 32-bit by
// construction (chunk_base is int32).
template <bool DIRECT, bool ASCII>
__global__ void __launch_bounds__(HASH_WARPS * 32, HASH_CTAS_PER_SM)
k_hash_chunks(const uint8_t *__restrict__ bytes, const int64_t *__restrict__ starts,
              const int32_t *__restrict__ lens, const int32_t *__restrict__ nchars,
              const int32_t *__restrict__ chunk_base, const int32_t *__restrict__ tile_first,
              const int32_t *__restrict__ chunk_off, int n, int cc, int64_t max_chunks,
              uint64_t *__restrict__ hashes, const unsigned long long *__restrict__ out_ptr)
{
    // out_ptr (partitioned multi-GPU path, see PartPlace): per request, where its hashes go --
    // an address in the owning GPU's inbound buffer; else they go to hashes[] (CSR order)
    __shared__ __align__(1024) uint8_t s_tile[HASH_WARPS * WTILE * 128];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int tile = blockIdx.x * HASH_WARPS + w;
    const int G0 = tile * WTILE;
    int total = chunk_base[n];
    if ((int64_t)total > max_chunks) total = (int)max_chunks;
    if (G0 >= total) return;                                   // whole warp

    // ---- map: which request owns each of the 32 chunks?
    ReqRec mine = { 0, 0, 0, 0, 0, 0 };   // record of the request that owns this lane's chunk
    const int r0 = tile_first[tile];
    for (int jb = 0;; jb += 32) {
        const int r = r0 + jb + lane;
        ReqRec q = { 0, 0, 0, 0, 0, 0 };
        bool live = false;
        if (r < n) {
            q.b = chunk_base[r]; q.b1 = chunk_base[r + 1];
            q.st = starts[r]; q.ln = lens[r];
            q.r = r;
            q.nc = (!ASCII && nchars) ? nchars[r] : q.ln;
            live = q.b < G0 + WTILE;
        }
        const bool nonempty = live && q.b1 > q.b;
        const int pos = q.b > G0 ? q.b - G0 : 0;
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
        const int tb = __shfl_sync(0xffffffffu, q.b, src);
        const int tb1 = ASCII ? 0 : __shfl_sync(0xffffffffu, q.b1, src);
        const int tln = __shfl_sync(0xffffffffu, q.ln, src);
        const int tnc = ASCII ? tln : __shfl_sync(0xffffffffu, q.nc, src);
        const long long tst = __shfl_sync(0xffffffffu, q.st, src);
        const int tr = DIRECT ? __shfl_sync(0xffffffffu, q.r, src) : 0;
        if (below) { mine.b = tb; mine.b1 = tb1; mine.ln = tln; mine.nc = tnc; mine.st = tst; mine.r = tr; }
        // more requests can start inside the tile only if the last lane's request ends inside it
        const int last_b1 = __shfl_sync(0xffffffffu, q.b1, 31);
        const bool last_live = __shfl_sync(0xffffffffu, (int)live, 31);
        if (!(last_live && last_b1 < G0 + WTILE && r0 + jb + 32 < n)) break;
    }

    // ---- chunk descriptor
    const int g = G0 + lane;
    const bool valid = g < total;
    const uint8_t *p = nullptr;
    int clen = 0, kind = 0;               // 0 none / done elsewhere, 1 full staged, 3 generic
    uint64_t *dst = nullptr;              // DIRECT: where this chunk's hash goes (owner's buffer)
    if (valid) {
        const int d = g - mine.b;         // depth of the chunk inside its request (0-based)
        if (DIRECT) dst = (uint64_t *)(uintptr_t)out_ptr[mine.r] + d;
        int off;
        bool last_ragged;                 // the request's ragged last chunk (ASCII requests only)
        if (ASCII || mine.nc == mine.ln) {               // ASCII: chunk = cc bytes
            off = d * cc;
            clen = min(cc, mine.ln - off);
            last_ragged = clen < cc;                     // only the last chunk can be short
        } else {                                         // code-point boundaries (k_utf8_bounds)
            off = chunk_off[g];
            const int end = (g + 1 < mine.b1 && (int64_t)g + 1 < max_chunks) ? chunk_off[g + 1] : mine.ln;
            clen = end - off;
            last_ragged = false;
        }
        p = bytes + mine.st + off;
        const bool aligned = (((uintptr_t)p & 15) == 0);
        const bool tail_done_by_prep = last_ragged && ((((uintptr_t)bytes + mine.st) & 15) == 0);
        kind = tail_done_by_prep ? 0 : ((aligned && clen == 128) ? 1 : 3);
    }

    // ---- stage this warp's full chunks, then hash
    uint64_t h = 0;
    const uint32_t wrow = (uint32_t)__cvta_generic_to_shared(s_tile) + (uint32_t)(w * WTILE) * 128;
    // staged chunks are 16-byte aligned: lanes exchange them as 32-bit offsets (in 16-byte
    // units, +1; 0 = not staged) from the buffer's aligned base
    const uintptr_t base16 = (uintptr_t)bytes & ~(uintptr_t)15;
    const uint32_t my_src = kind == 1 ? (uint32_t)(((uintptr_t)p - base16) >> 4) + 1u : 0u;
    // lane copies unit u of chunk c = 4k + sub; (c & 7) = sub ^ 4(k&1), so the swizzled
    // destination is one of two per-lane bases plus k * 512
    const int u = lane & 7, sub = lane >> 3;
    const uint64_t pol = l2_evict_first_policy();
    uint32_t d_even = wrow + sub * 128 + ((u ^ sub) << 4);
    asm volatile("" : "+r"(d_even));
    const uint32_t d_odd = d_even ^ 64u;
    const uint8_t *ubase = (const uint8_t *)base16 + (u << 4) - 16;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const uint32_t src = __shfl_sync(0xffffffffu, my_src, 4 * k + sub);
        if (src)
            cp_async16_stream(((k & 1) ? d_odd : d_even) + k * 512, ubase + ((size_t)src << 4), pol);
    }
    cp_async_wait_all();
    __syncwarp();
    if (kind == 1) h = xxh64_row128(wrow + lane * 128, lane & 7);
    else if (kind == 3) h = xxh64_global(p, clen);
    if (kind) {
        if (DIRECT) *dst = h;
        else hashes[g] = h;
    }
}

// ---------------------------------------------------------------------------
// k_hash_stream: the same work as k_hash_chunks with every warp as a PERSISTENT worker that
// keeps loads in flight all the time.  k_hash_chunks' warps live for one tile -- map (two
// dependent L2 round trips), stage, wait, hash -- so the loads of a warp are in flight for
// about a sixth of its life, and of 192 KB of staging memory per SM only ~32 KB carry data on
// average.  Here a warp owns two 4 KB buffers and walks tiles gw, gw + NW, gw + 2 NW, ...:
//     iteration for tile t:   issue the record loads of tile t+2 (its first request is known
//                             from a load issued one iteration earlier)
//                             wait for the staged bytes of tile t, hash, store
//                             finish the map of tile t+2 and stage it into the buffer just
//                             freed (tile t+1 has been in flight all along)
// so every dependent load has a whole hashing phase to arrive and one to two buffers per warp
// are always being filled.  3 CTAs x 8 warps per SM.
// ---------------------------------------------------------------------------
constexpr int HS_WARPS = 8, HS_CTAS_PER_SM = 3;
constexpr int HS_SMEM = HS_WARPS * 2 * WTILE * 128;          // 64 KB, dynamic

struct HsDesc { const uint8_t *p; uint64_t *dst; int clen, kind; };

__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait_group() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

template <bool DIRECT, bool ASCII>
__global__ void __launch_bounds__(HS_WARPS * 32, HS_CTAS_PER_SM)
k_hash_stream(const uint8_t *__restrict__ bytes, const int64_t *__restrict__ starts,
              const int32_t *__restrict__ lens, const int32_t *__restrict__ nchars,
              const int32_t *__restrict__ chunk_base, const int32_t *__restrict__ tile_first,
              const int32_t *__restrict__ chunk_off, int n, int cc, int64_t max_chunks,
              uint64_t *__restrict__ hashes, const unsigned long long *__restrict__ out_ptr)
{
    extern __shared__ __align__(1024) uint8_t hs_smem[];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int gw = blockIdx.x * HS_WARPS + w, NW = gridDim.x * HS_WARPS;
    int total = chunk_base[n];
    if ((int64_t)total > max_chunks) total = (int)max_chunks;
    const int ntiles = (total + WTILE - 1) / WTILE;
    if (gw >= ntiles) return;                                  // whole warp
    const uint32_t buf_base = (uint32_t)__cvta_generic_to_shared(hs_smem) + (uint32_t)w * (2 * WTILE * 128);
    uint32_t buf = buf_base;                                   // buffer of the tile hashed next
    const uintptr_t base16 = (uintptr_t)bytes & ~(uintptr_t)15;
    const int u = lane & 7, sub = lane >> 3;
    const uint64_t pol = l2_evict_first_policy();
    const uint8_t *ubase = (const uint8_t *)base16 + (u << 4) - 16;

    // B: the records of the first 32 requests a tile can touch (independent loads)
    auto load_recs = [&](int tile, int r0, bool &live) -> ReqRec {
        const int r = r0 + lane;
        ReqRec q = { 0, 0, 0, 0, 0, 0 };
        live = false;
        if (r < n) {
            q.b = chunk_base[r]; q.b1 = chunk_base[r + 1];
            q.st = starts[r]; q.ln = lens[r];
            q.r = r;
            q.nc = (!ASCII && nchars) ? nchars[r] : q.ln;
            live = q.b < tile * WTILE + WTILE;
        }
        return q;
    };
    // C: finish the map (each chunk lane pulls the record of its request), describe the chunk
    auto describe = [&](int tile, int r0, ReqRec q, bool live) -> HsDesc {
        const int G0 = tile * WTILE;
        ReqRec mine = { 0, 0, 0, 0, 0, 0 };
        for (int jb = 0;; jb += 32) {
            if (jb) q = load_recs(tile, r0 + jb, live);        // short requests: more than 32 per tile
            const bool nonempty = live && q.b1 > q.b;
            const int pos = q.b > G0 ? q.b - G0 : 0;
            const unsigned heads = __reduce_or_sync(0xffffffffu, nonempty ? (1u << pos) : 0u);
            const unsigned ne = __ballot_sync(0xffffffffu, nonempty);
            const unsigned below = heads & (0xffffffffu >> (31 - lane));
            int src = 0;
            if (below) {
                const int hp = 31 - __clz(below);
                const int k = __popc(heads & ((1u << hp) - 1));
                src = (ne == (0xffffffffu >> (32 - __popc(ne)))) ? k : (int)__fns(ne, 0, k + 1);
            }
            const int tb = __shfl_sync(0xffffffffu, q.b, src);
            const int tb1 = ASCII ? 0 : __shfl_sync(0xffffffffu, q.b1, src);
            const int tln = __shfl_sync(0xffffffffu, q.ln, src);
            const int tnc = ASCII ? tln : __shfl_sync(0xffffffffu, q.nc, src);
            const long long tst = __shfl_sync(0xffffffffu, q.st, src);
            const int tr = DIRECT ? __shfl_sync(0xffffffffu, q.r, src) : 0;
            if (below) { mine.b = tb; mine.b1 = tb1; mine.ln = tln; mine.nc = tnc; mine.st = tst; mine.r = tr; }
            const int last_b1 = __shfl_sync(0xffffffffu, q.b1, 31);
            const bool last_live = __shfl_sync(0xffffffffu, (int)live, 31);
            if (!(last_live && last_b1 < G0 + WTILE && r0 + jb + 32 < n)) break;
        }
        HsDesc d = { nullptr, nullptr, 0, 0 };
        const int g = G0 + lane;
        if (g < total) {
            const int dep = g - mine.b;
            d.dst = DIRECT ? (uint64_t *)(uintptr_t)out_ptr[mine.r] + dep : hashes + g;
            int off;
            bool last_ragged;
            if (ASCII || mine.nc == mine.ln) {
                off = dep * cc;
                d.clen = min(cc, mine.ln - off);
                last_ragged = d.clen < cc;
            } else {
                off = chunk_off[g];
                const int end = (g + 1 < mine.b1 && (int64_t)g + 1 < max_chunks) ? chunk_off[g + 1] : mine.ln;
                d.clen = end - off;
                last_ragged = false;
            }
            d.p = bytes + mine.st + off;
            const bool aligned = (((uintptr_t)d.p & 15) == 0);
            const bool tail_done_by_prep = last_ragged && ((((uintptr_t)bytes + mine.st) & 15) == 0);
            d.kind = tail_done_by_prep ? 0 : ((aligned && d.clen == 128) ? 1 : 3);
        }
        return d;
    };
    // stage the tile's full chunks into `b` (one commit group per tile, empty or not)
    auto stage = [&](const HsDesc &d, uint32_t b) {
        const uint32_t my_src = d.kind == 1 ? (uint32_t)(((uintptr_t)d.p - base16) >> 4) + 1u : 0u;
        uint32_t d_even = b + sub * 128 + ((u ^ sub) << 4);
        const uint32_t d_odd = d_even ^ 64u;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const uint32_t src = __shfl_sync(0xffffffffu, my_src, 4 * k + sub);
            if (src)
                cp_async16_stream(((k & 1) ? d_odd : d_even) + k * 512, ubase + ((size_t)src << 4), pol);
        }
        cp_async_commit();
    };
    auto hash_tile = [&](const HsDesc &d, uint32_t b) {
        uint64_t h = 0;
        if (d.kind == 1) h = xxh64_row128(b + lane * 128, lane & 7);
        else if (d.kind == 3) h = xxh64_global(d.p, d.clen);
        if (d.kind) *d.dst = h;
    };

    // ---- prologue: tiles t and t + NW staged, first request of t + 2 NW on its way
    int t = gw;
    bool live;
    HsDesc d0, d1 = { nullptr, nullptr, 0, 0 };
    {
        const int r0 = tile_first[t];
        const bool h1 = t + NW < ntiles;
        const int r1 = h1 ? tile_first[t + NW] : 0;
        ReqRec q = load_recs(t, r0, live);
        d0 = describe(t, r0, q, live);
        stage(d0, buf);
        if (h1) {
            q = load_recs(t + NW, r1, live);
            d1 = describe(t + NW, r1, q, live);
            stage(d1, buf_base + WTILE * 128);
        }
    }
    bool has1 = t + NW < ntiles;
    int r0_next = (t + 2 * NW < ntiles) ? tile_first[t + 2 * NW] : 0;
    for (;;) {
        const bool has2 = t + 2 * NW < ntiles;
        ReqRec q2 = { 0, 0, 0, 0, 0, 0 };
        bool live2 = false;
        if (has2) q2 = load_recs(t + 2 * NW, r0_next, live2);
        const int r0_next3 = (t + 3 * NW < ntiles) ? tile_first[t + 3 * NW] : 0;
        if (has1) cp_async_wait_group<1>(); else cp_async_wait_group<0>();
        __syncwarp();
        hash_tile(d0, buf);
        __syncwarp();                                         // the buffer is free again
        HsDesc d2 = { nullptr, nullptr, 0, 0 };
        if (has2) {
            d2 = describe(t + 2 * NW, r0_next, q2, live2);
            stage(d2, buf);
        }
        if (!has1) break;
        d0 = d1; d1 = d2; has1 = has2;
        t += NW;
        buf = (buf == buf_base) ? buf_base + WTILE * 128 : buf_base;
        r0_next = r0_next3;
    }
}

// =========================================================================
// Route stage.  a5 step 1 is k_walk<W_CLAIM> (trie.cuh): every node of every path is
// found or created and classified against the pre-batch masks
//   slots[g] = log index | PREOK_BIT   node usable before the batch (mask & avail != 0)
//   slots[g] = log index               otherwise; rec.first = (epoch, ~min request index)
//   dpre[i]  = number of leading PREOK nodes of request i (masks are nested along a path:
//              inserts add the endpoint to every node of it, so PREOK is a prefix property)
//   cand[i]  = mask & avail of the deepest PREOK node (avail when there is none)
// =========================================================================

// a5 step 2: deepest node usable by request i = last depth d with
//   PREOK(node) or first(node) < i      (monotone in d)
// If that node was usable before the batch (or depth 0) the request is a root
// and draws from its candidate set; else it follows the request that touched
// the node first.  The walk starts at dpre[i]: everything above is PREOK.
__device__ __forceinline__ void
select_request(int i, const Rec *__restrict__ log, const int32_t *__restrict__ chunk_base,
               const int32_t *__restrict__ chunk_end, int64_t max_chunks, const uint32_t *__restrict__ slots,
               const int32_t *__restrict__ dpre, const uint64_t *__restrict__ cand_in,
               unsigned long long tag, uint64_t seed, uint64_t seq0, int32_t *__restrict__ leader,
               int32_t *__restrict__ choice, int32_t *__restrict__ depth_out,
               uint64_t *__restrict__ matched_out, uint32_t req_off, const uint32_t *__restrict__ pos)
{
    const ReqSpan sp = req_span(chunk_base, chunk_end, i, max_chunks);
    const uint32_t *sl = slots + sp.b0;
    const int d0 = dpre[i];
    // "node d was touched by an earlier request of this batch" is a prefix property (whoever
    // touched a node touched everything above it on the same path, no later): find the last
    // such depth -- linear for the first few, then by bisection (long prompts: C4 has 256)
    const uint32_t me = pos ? pos[i] : req_off + (uint32_t)i;
    auto touched = [&](int d, uint32_t &f) -> bool {
        const uint32_t x = sl[d];
        if (x == SLOT_NONE) return false;
        const unsigned long long fv = log[x & IDX_MASK].first;
        f = ~(uint32_t)fv;
        return ((fv >> 32) == (tag >> 32)) && f < me;
    };
    int dstar = d0;
    uint32_t star_first = 0;
    int lo = d0, probes = 0;
    while (lo < sp.C && probes < 4) {
        uint32_t f;
        if (!touched(lo, f)) break;
        star_first = f; dstar = ++lo; probes++;
    }
    if (probes == 4 && lo < sp.C) {
        int hi = sp.C;                          // invariant: depths < lo touched, depth hi not
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            uint32_t f;
            if (touched(mid, f)) { star_first = f; dstar = mid + 1; lo = mid + 1; }
            else hi = mid;
        }
    }
    depth_out[i] = dstar;
    if (dstar == d0) {
        const uint64_t cand = cand_in[i];
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
k_select(const Rec *__restrict__ log, const int32_t *__restrict__ chunk_base,
         const int32_t *__restrict__ chunk_end, int n, int64_t max_chunks,
         const uint32_t *__restrict__ slots, const int32_t *__restrict__ dpre,
         const uint64_t *__restrict__ cand, unsigned long long tag, uint64_t seed,
         uint64_t seq0, int32_t *__restrict__ leader, int32_t *__restrict__ choice,
         int32_t *__restrict__ depth_out, uint64_t *__restrict__ matched_out, uint32_t req_off,
         const P2PDyn *__restrict__ dyn, const uint32_t *__restrict__ pos,
         const int32_t *__restrict__ n_dev)
{
    // one thread per request.  req_off = position of request 0 in the global order (0 on a
    // single GPU).  With `dyn` (partitioned multi-GPU path) seq0 lives on the device and pos[i]
    // is the request's global position, used for the tie-break counter only.
    if (n_dev) n = *n_dev;
    if (dyn) { req_off = dyn->my_off; seq0 = dyn->seq_base; }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        select_request(i, log, chunk_base, chunk_end, max_chunks, slots, dpre, cand, tag, seed, seq0, leader,
                       choice, depth_out, matched_out, req_off, pos);
}

// a5 step 3: followers take their (transitive) leader's endpoint; every
// request ORs its endpoint into the nodes below its match depth -- exactly the
// nodes it was the first to touch, so no other request writes them.
__device__ __forceinline__ void
commit_request(int i, Rec *__restrict__ log, TrieDev *dev, const int32_t *__restrict__ chunk_base,
               const int32_t *__restrict__ chunk_end, int64_t max_chunks, const uint32_t *__restrict__ slots,
               const int32_t *__restrict__ leader, const int32_t *__restrict__ choice,
               const int32_t *__restrict__ depth, int32_t *__restrict__ endpoint,
               uint64_t *__restrict__ matched_out, const uint32_t *__restrict__ idx_of_pos)
{
    // leaders are global positions; idx_of_pos (partitioned multi-GPU path, where requests sit
    // in arrival order, not in global order) maps them back to local indices
    int j = i, l;
    while ((l = leader[j]) >= 0) j = idx_of_pos ? (int)idx_of_pos[l] : l;   // positions strictly decrease
    const int e = choice[j];
    endpoint[i] = e;
    if (j != i) matched_out[i] = (e >= 0) ? (1ULL << e) : 0;
    if (e < 0) return;
    const uint64_t bit = 1ULL << e;
    if (!(dev->root_mask & bit)) atomicOr((unsigned long long *)&dev->root_mask, bit);
    const ReqSpan sp = req_span(chunk_base, chunk_end, i, max_chunks);
    for (int d = depth[i]; d < sp.C; d++) {
        const uint32_t x = slots[sp.b0 + d];
        if (x != SLOT_NONE) atomicOr((unsigned long long *)&log[x & IDX_MASK].mask, bit);
    }
}

__global__ void __launch_bounds__(256)
k_commit(Rec *__restrict__ log, TrieDev *dev, const int32_t *__restrict__ chunk_base,
         const int32_t *__restrict__ chunk_end, int n, int64_t max_chunks, const uint32_t *__restrict__ slots, const int32_t *__restrict__ leader,
         const int32_t *__restrict__ choice, const int32_t *__restrict__ depth,
         int32_t *__restrict__ endpoint, uint64_t *__restrict__ matched_out,
         const int32_t *__restrict__ n_dev, const uint32_t *__restrict__ idx_of_pos)
{
    if (n_dev) n = *n_dev;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        commit_request(i, log, dev, chunk_base, chunk_end, max_chunks, slots, leader, choice, depth, endpoint,
                       matched_out, idx_of_pos);
}


// select + commit in ONE launch (single GPU, short prompts): a request publishes, in one tagged
// 64-bit word, either its endpoint (roots) or the position of its leader (followers); followers
// then chase their chain through those words, waiting for words that are not there yet.  Leaders
// have strictly smaller positions, i.e. live in the same or an earlier CTA, and nobody waits
// before it has published, so the waits always end.  Saves a launch and a ramp on the route
// step's dependent chain.
__global__ void __launch_bounds__(256)
k_select_commit(Rec *__restrict__ log, TrieDev *dev, const int32_t *__restrict__ chunk_base,
                int n, int64_t max_chunks, const uint32_t *__restrict__ slots,
                const int32_t *__restrict__ dpre, const uint64_t *__restrict__ cand,
                unsigned long long tag, uint64_t seed, uint64_t seq0, int32_t *__restrict__ leader,
                int32_t *__restrict__ choice, int32_t *__restrict__ depth_out,
                uint64_t *__restrict__ matched_out, int32_t *__restrict__ endpoint,
                unsigned long long *res)
{
    constexpr unsigned long long ROOT = 1ULL << 31;
    for (int base = blockIdx.x * blockDim.x; base < n; base += gridDim.x * blockDim.x) {
        const int i = base + threadIdx.x;
        const bool active = i < n;
        int l = -1, ch = -1;
        if (active) {
            select_request(i, log, chunk_base, nullptr, max_chunks, slots, dpre, cand, tag, seed, seq0, leader,
                           choice, depth_out, matched_out, 0u, nullptr);
            l = leader[i]; ch = choice[i];
            const unsigned long long wv = tag | (l < 0 ? (ROOT | (unsigned long long)(unsigned)(ch + 1))
                                                      : (unsigned long long)(unsigned)l);
            *(volatile unsigned long long *)&res[i] = wv;
        }
        __syncwarp();
        if (!active) continue;
        int e = ch;
        if (l >= 0) {
            int j = l;
            for (;;) {
                unsigned long long v;
                do { v = *(volatile unsigned long long *)&res[j]; } while ((v >> 32) != (tag >> 32));
                if (v & ROOT) { e = (int)(v & 0xFFu) - 1; break; }
                j = (int)(v & 0x7FFFFFFFu);
            }
            matched_out[i] = (e >= 0) ? (1ULL << e) : 0;
        }
        endpoint[i] = e;
        if (e < 0) continue;
        const uint64_t bit = 1ULL << e;
        if (!(dev->root_mask & bit)) atomicOr((unsigned long long *)&dev->root_mask, bit);
        const ReqSpan sp = req_span(chunk_base, nullptr, i, max_chunks);
        for (int d = depth_out[i]; d < sp.C; d++) {
            const uint32_t x = slots[sp.b0 + d];
            if (x != SLOT_NONE) atomicOr((unsigned long long *)&log[x & IDX_MASK].mask, bit);
        }
    }
}

// The same with a warp per request, lanes over depths: long prompts (C4: 256 chunks, half of
// them new) make the thread version a chain of a hundred dependent atomics.
__global__ void __launch_bounds__(256)
k_commit_warp(Rec *__restrict__ log, TrieDev *dev, const int32_t *__restrict__ chunk_base,
              const int32_t *__restrict__ chunk_end, int n, int64_t max_chunks, const uint32_t *__restrict__ slots, const int32_t *__restrict__ leader,
              const int32_t *__restrict__ choice, const int32_t *__restrict__ depth,
              int32_t *__restrict__ endpoint, uint64_t *__restrict__ matched_out,
              const int32_t *__restrict__ n_dev, const uint32_t *__restrict__ idx_of_pos)
{
    if (n_dev) n = *n_dev;
    const int lane = threadIdx.x & 31;
    const int w0 = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    for (int i = w0; i < n; i += nw) {
        int e = -1;
        if (lane == 0) {
            int j = i, l;
            while ((l = leader[j]) >= 0) j = idx_of_pos ? (int)idx_of_pos[l] : l;
            e = choice[j];
            endpoint[i] = e;
            if (j != i) matched_out[i] = (e >= 0) ? (1ULL << e) : 0;
            if (e >= 0 && !(dev->root_mask & (1ULL << e)))
                atomicOr((unsigned long long *)&dev->root_mask, 1ULL << e);
        }
        e = __shfl_sync(0xffffffffu, e, 0);
        if (e < 0) continue;
        const uint64_t bit = 1ULL << e;
        const ReqSpan sp = req_span(chunk_base, chunk_end, i, max_chunks);
        for (int d = depth[i] + lane; d < sp.C; d += 32) {
            const uint32_t x = slots[sp.b0 + d];
            if (x != SLOT_NONE) atomicOr((unsigned long long *)&log[x & IDX_MASK].mask, bit);
        }
    }
}

// =========================================================================
// Latency path: ONE request per launch (live traffic arrives one request at a time and a
// batch of one through the batch kernels costs six launches and two copies).  One CTA:
//   1. the prompt is read straight from the caller's pinned host buffer into shared memory
//      (16-byte loads, all in flight at once: one PCIe round trip), hashed from there, a
//      thread per chunk;
//   2. warp 0 walks the trie with the batch path's own walker (claim mode, 32 lanes);
//   3. a lone request is a root by definition: draw from the candidates of the deepest node
//      usable before the call, OR the endpoint bit into everything below, write the result to
//      the caller's (host-mapped) result block.
// ASCII prompts only (code points = bytes); anything else is reported back untouched and the
// caller takes the batch path.
// =========================================================================
constexpr int ONE_PASS = 32768;        // prompt bytes staged per pass
constexpr int ONE_MAX_CHUNKS = 16384;  // scratch capacity of the latency path
constexpr uint32_t ONE_DONE = 0x80000000u, ONE_NONASCII = 0x40000000u;
struct OneResult { int32_t endpoint, depth; uint64_t matched; uint32_t status, pad[3]; };

__device__ __forceinline__ uint64_t xxh64_words(const uint8_t *p, int len)
{   // p 8-byte aligned (shared memory)
    const uint8_t *end = p + len;
    uint64_t h;
    if (len >= 32) {
        uint64_t v1 = XP1 + XP2, v2 = XP2, v3 = 0, v4 = 0 - XP1;
        const uint8_t *lim = end - 32;
        do {
            const uint4 a = *(const uint4 *)p, b = *(const uint4 *)(p + 16);
            v1 = xround(v1, u64_of(a.x, a.y));
            v2 = xround(v2, u64_of(a.z, a.w));
            v3 = xround(v3, u64_of(b.x, b.y));
            v4 = xround(v4, u64_of(b.z, b.w));
            p += 32;
        } while (p <= lim);
        h = rotl<1>(v1) + rotl<7>(v2) + rotl<12>(v3) + rotl<18>(v4);
        h = xmerge(h, v1); h = xmerge(h, v2); h = xmerge(h, v3); h = xmerge(h, v4);
    } else {
        h = XP5;
    }
    h += (uint64_t)len;
    while (p + 8 <= end) { h ^= xround(0, *(const uint64_t *)p); h = mulc<XP1>(rotl<27>(h)) + XP4; p += 8; }
    if (p + 4 <= end) { h ^= mulc<XP1>((uint64_t)(*(const uint32_t *)p)); h = mulc<XP2>(rotl<23>(h)) + XP3; p += 4; }
    while (p < end) { h ^= mulc<XP5>((uint64_t)(*p)); h = mulc<XP1>(rotl<11>(h)); p++; }
    return xavalanche(h);
}

__global__ void __launch_bounds__(256)
k_route_one(WalkArgs a, TrieDev *dev, const uint8_t *__restrict__ prompt, int len, int cc,
            uint64_t seed, uint64_t seq, int32_t *chunk_base, uint64_t *hashes, OneResult *out)
{
    __shared__ __align__(16) uint8_t s_buf[ONE_PASS];
    __shared__ unsigned s_hi;
    const int C = (len + cc - 1) / cc;
    const int pass = (ONE_PASS / cc) * cc;             // whole chunks per pass; cc % 16 == 0
    if (threadIdx.x == 0) s_hi = 0;
    unsigned hi = 0;
    for (int base = 0; base < len; base += pass) {
        const int m = min(pass, len - base);
        __syncthreads();
#pragma unroll 4
        for (int x = threadIdx.x; x * 16 < m; x += blockDim.x) {
            const uint4 v = *(const uint4 *)(prompt + base + x * 16);
            ((uint4 *)s_buf)[x] = v;
            hi |= (v.x | v.y | v.z | v.w) & 0x80808080u;
        }
        __syncthreads();
        for (int c = threadIdx.x; c * cc < m; c += blockDim.x)
            hashes[base / cc + c] = xxh64_words(s_buf + c * cc, min(cc, m - c * cc));
    }
    if (hi) atomicOr(&s_hi, 1u);
    if (threadIdx.x == 0) { chunk_base[0] = 0; chunk_base[1] = C; }
    __syncthreads();
    if (threadIdx.x >= 32) return;
    const int lane = threadIdx.x;
    if (s_hi) {                                         // not ASCII: nothing touched
        if (lane == 0) { out->endpoint = -1; out->depth = 0; out->matched = 0; __threadfence_system(); out->status = ONE_DONE | ONE_NONASCII; }
        return;
    }
    const Tile g = make_tile<32>(lane);
    WarpAlloc wa;
    { const uint2 v = a.t.wblk[NWSLOT - 1]; wa.cur = v.x; wa.end = v.y; }
    ReqSpan sp; sp.b0 = 0; sp.C = C;
    Pending none = { -1, 0, 0, 0 };
    walk_request<W_CLAIM, false, 32>(a, 0, g, wa, sp, lane < C ? __ldcg(hashes + lane) : 0, none);
    if (lane == 0) a.t.wblk[NWSLOT - 1] = make_uint2(wa.cur, wa.end);
    __syncwarp();
    const int dpre = __ldcg(a.dpre);
    const uint64_t cand = __ldcg(a.cand);
    const int cnt = __popcll(cand);
    const int e = cnt ? nth_set_bit(cand, pick(seed, seq, (uint32_t)cnt)) : -1;
    if (e >= 0) {
        const uint64_t bit = 1ULL << e;
        if (lane == 0 && !(dev->root_mask & bit)) atomicOr((unsigned long long *)&dev->root_mask, bit);
        for (int d = dpre + lane; d < C; d += 32) {
            const uint32_t x = __ldcg(a.slots + d);
            if (x != SLOT_NONE) atomicOr((unsigned long long *)&a.t.log[x & IDX_MASK].mask, bit);
        }
    }
    if (lane == 0) {
        out->endpoint = e; out->depth = dpre; out->matched = cand;
        __threadfence_system();
        out->status = ONE_DONE | (dev->status & 0xFFFFu);
    }
}

// ------------------------------------------------------------ dump / restore / sweeps
__global__ void k_count_nodes(const Head *__restrict__ tab, const Rec *__restrict__ log, int64_t cap,
                              TrieDev *dev)
{
    unsigned long long mine = 0;
    for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < cap;
         s += (int64_t)gridDim.x * blockDim.x) {
        const Head e = tab[s];
        if (e.pk) mine += (uint32_t)log[(uint32_t)e.pk].rc;
    }
    for (int o = 16; o; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
    if ((threadIdx.x & 31) == 0 && mine) atomicAdd(&dev->num_nodes, mine);
}

// One thread per run head: the run's records in depth order, as
//   {log index | records left in the run << 32, parent log index (ROOT_IDX for depth-1 nodes)
//    | 1 << 63 on the first node of a run, chunk hash, mask}
__global__ void k_dump(const Head *__restrict__ tab, const Rec *__restrict__ log, int64_t cap,
                       TrieDev *dev, uint64_t *__restrict__ out, int64_t max_nodes)
{
    for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < cap;
         s += (int64_t)gridDim.x * blockDim.x) {
        const Head e = tab[s];
        if (!e.pk) continue;
        const uint32_t idx = (uint32_t)e.pk;
        const uint32_t len = (uint32_t)log[idx].rc;
        const unsigned long long at = atomicAdd(&dev->dump_count, (unsigned long long)len);
        for (uint32_t j = 0; j < len; j++) {
            if ((int64_t)(at + j) >= max_nodes) break;
            const Rec r = log[idx + j];
            out[4 * (at + j) + 0] = (uint64_t)(idx + j) | ((uint64_t)(len - j) << 32);
            out[4 * (at + j) + 1] = j ? (uint64_t)(idx + j - 1) : ((uint64_t)((uint32_t)(e.pk >> 32) - 2u) | (1ULL << 63));
            out[4 * (at + j) + 2] = r.h;
            out[4 * (at + j) + 3] = r.mask;
        }
    }
}

// Inverse of k_dump into an EMPTY table: records go back to their log indices, run heads
// back into the hash table, so the restored table has the layout (and speed) of the saved one.
__global__ void k_restore(TrieRef t, const uint64_t *__restrict__ ent, int64_t count, uint64_t root_or)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0 && root_or) atomicOr((unsigned long long *)&t.dev->root_mask, root_or);
    if (i >= count) return;
    const uint64_t e0 = ent[4 * i], e1 = ent[4 * i + 1], h = ent[4 * i + 2], m = ent[4 * i + 3];
    const uint32_t idx = (uint32_t)e0, rem = (uint32_t)(e0 >> 32), parent = (uint32_t)e1;
    if (rem == 0 || (uint64_t)idx + rem > t.log_cap) {
        atomicOr(&t.dev->status, B200R_S_TABLE_FULL);
        return;
    }
    st_rec_key(t.log + idx, h, rem);
    t.log[idx].mask = m;
    atomicMax(&t.dev->log_count, (unsigned long long)idx + 1);
    if ((e1 >> 63) && head_publish(t.tab, t.tab_log2, h, parent_key(parent), idx) != idx)
        atomicOr(&t.dev->status, B200R_S_INTERNAL);
}

// epoch wrap: forget every first-toucher word (masks and hashes stay)
__global__ void k_sweep_first(Rec *__restrict__ log, const TrieDev *__restrict__ dev, uint32_t log_cap)
{
    unsigned long long n = dev->log_count;
    if (n > log_cap) n = log_cap;
    for (unsigned long long s = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; s < n;
         s += (unsigned long long)gridDim.x * blockDim.x)
        log[s].first = 0;
}

// endpoints retired for good: take their bits out of every endpoint set
__global__ void k_clear_bits(Rec *__restrict__ log, TrieDev *dev, uint32_t log_cap, uint64_t bits)
{
    unsigned long long n = dev->log_count;
    if (n > log_cap) n = log_cap;
    if (blockIdx.x == 0 && threadIdx.x == 0) dev->root_mask &= ~bits;
    for (unsigned long long s = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; s < n;
         s += (unsigned long long)gridDim.x * blockDim.x) {
        const uint64_t m = log[s].mask;
        if (m & bits) log[s].mask = m & ~bits;
    }
}

}  // namespace b200r

// =========================================================================
// Host side shared with multigpu.cu
// =========================================================================
namespace b200r {

Plan make_plan(int64_t n, int64_t max_chunks, uint32_t flags, bool routing)
{
    Plan p;
    size_t o = 0;
    const int64_t ntiles = (max_chunks + 31) / 32 + 1;
    const int64_t nblk = (n + PREP_BLOCK - 1) / PREP_BLOCK + 1;
    const bool ascii = flags & B200R_F_ALL_ASCII;
    p.off_tile_first = o; o += align_up(ntiles * 4);
    p.off_block_sums = o; o += align_up(nblk * 4);
    p.off_nchars = o;     o += ascii ? 0 : align_up((n + 1) * 4);
    p.off_chunk_off = o;  o += ascii ? 0 : align_up((max_chunks + 1) * 4);
    p.off_chunk_base = o; o += routing ? align_up((n + 1) * 4) : 0;
    p.off_hashes = o;     o += routing ? align_up((max_chunks + 1) * 8) : 0;
    p.off_slots = o;      o += routing ? align_up((max_chunks + 1) * 4) : 0;
    p.off_leader = o;     o += routing ? align_up((n + 1) * 4) : 0;
    p.off_choice = o;     o += routing ? align_up((n + 1) * 4) : 0;
    p.off_dpre = o;       o += routing ? align_up((n + 1) * 4) : 0;
    p.off_cand = o;       o += routing ? align_up((n + 1) * 8) : 0;
    p.off_res = o;        o += routing ? align_up((n + 1) * 8) : 0;
    p.total = o + 256;
    return p;
}

int check_prompt_args(const void *bytes, const void *starts, const void *lens, int64_t n,
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

// prep + (utf8) + hash; shared by b200r_xxh64_chunks and the routing entry points
int launch_hash(const uint8_t *d_bytes, const int64_t *d_starts, const int32_t *d_lens,
                const int32_t *d_nchars, int64_t n, int32_t cc, uint32_t flags,
                int32_t *d_chunk_base, uint64_t *d_hashes, int64_t max_chunks,
                uint8_t *scratch, const Plan &p, TrieDev *dev, cudaStream_t st, int phase,
                const PartPlace *pp)
{
    PartPlace place = PartPlace();          // world == 0: off
    if (pp) place = *pp;
    // phase 0: everything; 1: all but the chunk-hashing kernel (the "prep" part: code-point
    // counts, CSR, tile map, UTF-8 chunk offsets, ragged tails); 2: the chunk-hashing kernel only
    int32_t *tile_first = (int32_t *)(scratch + p.off_tile_first);
    int32_t *block_sums = (int32_t *)(scratch + p.off_block_sums);
    int32_t *chunk_off = nullptr;
    const bool ascii = flags & B200R_F_ALL_ASCII;
    const int nwarp_blocks = (int)((n + WARPS_PER_CTA - 1) / WARPS_PER_CTA);
    if (ascii) {
        d_nchars = nullptr;                     // nchars == lens by contract
    } else {
        chunk_off = (int32_t *)(scratch + p.off_chunk_off);
        if (!d_nchars && phase == 2) d_nchars = (int32_t *)(scratch + p.off_nchars);
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
    if (phase != 2) {
    { ProfScope ps__(PK_PREP, st); g_launches += 1;
    k_prep_sums<<<nblk, PREP_BLOCK, 0, st>>>(d_lens, d_nchars, (int)n, cc, block_sums, nullptr);
    B200R_LAUNCH_CHECK("k_prep_sums");
    }
    { ProfScope ps__(PK_PREP, st); g_launches += 1;
    k_prep_scan<<<nblk, PREP_BLOCK, 0, st>>>(d_lens, d_nchars, (int)n, cc, block_sums, d_chunk_base,
                                             tile_first, max_chunks, dev, d_bytes, d_starts, d_hashes, nullptr,
                                             place);
    B200R_LAUNCH_CHECK("k_prep_scan");
    }
    if (!ascii) {
        { ProfScope ps__(PK_UTF8, st); g_launches += 1;
        k_utf8_bounds<<<nwarp_blocks, WARPS_PER_CTA * 32, 0, st>>>(
            d_bytes, d_starts, d_lens, d_nchars, d_chunk_base, (int)n, cc, max_chunks, chunk_off);
        B200R_LAUNCH_CHECK("k_utf8_bounds");
        }
    }
    }
    const int ntiles = (int)((max_chunks + HASH_WARPS * WTILE - 1) / (HASH_WARPS * WTILE));
    if (ntiles > 0 && phase != 1) {
        ProfScope ps__(PK_HASH, st); g_launches += 1;
        const unsigned long long *op = pp ? pp->out_ptr : nullptr;
#define B200R_HASH_ARGS d_bytes, d_starts, d_lens, d_nchars, d_chunk_base, tile_first, chunk_off, (int)n, cc, \
                        max_chunks, d_hashes, op
        if (g_hash_mode == 1) {
            if (pp)            k_hash_chunks<true, true><<<ntiles, HASH_WARPS * 32, 0, st>>>(B200R_HASH_ARGS);
            else if (ascii)    k_hash_chunks<false, true><<<ntiles, HASH_WARPS * 32, 0, st>>>(B200R_HASH_ARGS);
            else               k_hash_chunks<false, false><<<ntiles, HASH_WARPS * 32, 0, st>>>(B200R_HASH_ARGS);
        } else {
            static bool attr_set = false;
            if (!attr_set) {
                B200R_CUDA(cudaFuncSetAttribute(k_hash_stream<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, HS_SMEM));
                B200R_CUDA(cudaFuncSetAttribute(k_hash_stream<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, HS_SMEM));
                B200R_CUDA(cudaFuncSetAttribute(k_hash_stream<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, HS_SMEM));
                attr_set = true;
            }
            int grid = ntiles < 148 * HS_CTAS_PER_SM ? ntiles : 148 * HS_CTAS_PER_SM;
            if (pp)            k_hash_stream<true, true><<<grid, HS_WARPS * 32, HS_SMEM, st>>>(B200R_HASH_ARGS);
            else if (ascii)    k_hash_stream<false, true><<<grid, HS_WARPS * 32, HS_SMEM, st>>>(B200R_HASH_ARGS);
            else               k_hash_stream<false, false><<<grid, HS_WARPS * 32, HS_SMEM, st>>>(B200R_HASH_ARGS);
        }
#undef B200R_HASH_ARGS
        B200R_LAUNCH_CHECK("k_hash_chunks");
    }
    return B200R_OK;
}

TrieRef trie_ref(const b200r_trie *t)
{
    TrieRef r;
    r.tab = t->tab; r.log = t->log; r.dev = t->dev; r.wblk = t->wblk;
    r.tab_log2 = t->cap_log2 + 1;
    r.log_cap = (uint32_t)1 << t->cap_log2;
    return r;
}

// Batch epoch of the "first toucher" words: value = (epoch << 32) | ~request, merged with
// atomicMax, so a newer batch overrides whatever older batches left and, inside a batch, the
// smallest request index wins.  Zero (fresh table) means "never".  When the 32-bit epoch
// runs out, one sweep zeroes the words and counting restarts at 1; hashes and masks are
// untouched, so results do not change.
static int next_epoch(b200r_trie *t, cudaStream_t st, unsigned long long *tag)
{
    if (t->epoch >= 0xFFFFFFF0u) {
        g_launches += 1;
        k_sweep_first<<<148 * 8, 256, 0, st>>>(t->log, t->dev, (uint32_t)1 << t->cap_log2);
        B200R_LAUNCH_CHECK("k_sweep_first");
        t->epoch = 0;
    }
    t->epoch++;
    *tag = (unsigned long long)t->epoch << 32;
    return B200R_OK;
}

// How the claim step is decomposed: 1 = a thread per request, 2 = a warp, 3 = 16 lanes, 4 = 8
// lanes (identical results).  Measured on C3 (21.5 chunks per prompt): thread 30 us, 8 lanes 39,
// 16 lanes 47, warp 49.5 per 64K batch -- short prompts leave most lanes of a wider tile idle
// while costing its issue slots all the same.  Long prompts (C4: 256 chunks) want the coalesced
// record reads of the warp.
int claim_kind_for(int64_t n, int64_t chunks)
{
    if (g_claim_mode) return g_claim_mode;
    return chunks <= n * 64 ? 1 : 2;
}

static int walk_grid(int64_t n, int tw = 32)
{
    int64_t b = (n * tw + WARPS_PER_CTA * 32 - 1) / (WARPS_PER_CTA * 32);
    if (b > WALK_MAX_CTAS) b = WALK_MAX_CTAS;
    return b < 1 ? 1 : (int)b;
}

// claim -> select -> commit; advances the table's batch epoch
int launch_route_core(b200r_trie *t, const RouteCore &rc, cudaStream_t st)
{
    unsigned long long tag;
    int rc0 = next_epoch(t, st, &tag);
    if (rc0) return rc0;
    WalkArgs a = WalkArgs();
    a.t = trie_ref(t);
    a.hashes = rc.hashes; a.chunk_base = rc.chunk_base; a.chunk_end = rc.chunk_end;
    a.max_chunks = rc.max_chunks;
    a.avail = rc.avail; a.tag = tag; a.slots = rc.slots; a.dpre = rc.dpre; a.cand = rc.cand;
    a.pos = rc.pos;
    const bool fused_sc = rc.res && !rc.pos && !rc.idx_of_pos && !rc.dyn && !rc.n_dev && !rc.chunk_end &&
                          rc.claim_kind == 1 && g_fuse_sc;
    a.res = fused_sc ? rc.res : nullptr;
    const int nn = rc.n_dev ? 0 : (int)rc.n;
    const int tblocks = (int)((rc.n + 255) / 256);
    { ProfScope ps__(PK_CLAIM, st); g_launches += 1;
    if (rc.claim_kind == 1) {
        int64_t blocks = (rc.n + 127) / 128;
        if (blocks > 148 * 16) blocks = 148 * 16;
        k_claim_thread<<<(int)(blocks < 1 ? 1 : blocks), 128, 0, st>>>(a, nn, rc.n_dev);
    } else if (rc.claim_kind == 4) {
        k_walk<W_CLAIM, 8><<<walk_grid(rc.n, 8), WARPS_PER_CTA * 32, 0, st>>>(a, nn, rc.n_dev);
    } else if (rc.claim_kind == 3) {
        k_walk<W_CLAIM, 16><<<walk_grid(rc.n, 16), WARPS_PER_CTA * 32, 0, st>>>(a, nn, rc.n_dev);
    } else {
        k_walk<W_CLAIM, 32><<<walk_grid(rc.n), WARPS_PER_CTA * 32, 0, st>>>(a, nn, rc.n_dev);
    }
    B200R_LAUNCH_CHECK("k_walk<claim>");
    }
    if (fused_sc) {
        ProfScope ps__(PK_SELECT, st); g_launches += 1;
        k_select_commit<<<tblocks, 256, 0, st>>>(t->log, t->dev, rc.chunk_base, nn, rc.max_chunks, rc.slots, rc.dpre,
                                                rc.cand, tag, rc.seed, rc.seq0, rc.leader, rc.choice, rc.depth,
                                                rc.matched, rc.endpoint, rc.res);
        B200R_LAUNCH_CHECK("k_select_commit");
        return B200R_OK;
    }
    { ProfScope ps__(PK_SELECT, st); g_launches += 1;
    k_select<<<tblocks, 256, 0, st>>>(t->log, rc.chunk_base, rc.chunk_end, nn, rc.max_chunks, rc.slots, rc.dpre,
                                     rc.cand, tag, rc.seed, rc.seq0, rc.leader, rc.choice,
                                     rc.depth, rc.matched, 0u, rc.dyn, rc.pos, rc.n_dev);
    B200R_LAUNCH_CHECK("k_select");
    }
    { ProfScope ps__(PK_COMMIT, st); g_launches += 1;
    if (rc.claim_kind == 1)
        k_commit<<<tblocks, 256, 0, st>>>(t->log, t->dev, rc.chunk_base, rc.chunk_end, nn, rc.max_chunks, rc.slots,
                                         rc.leader, rc.choice, rc.depth, rc.endpoint, rc.matched, rc.n_dev,
                                         rc.idx_of_pos);
    else
        k_commit_warp<<<walk_grid(rc.n), 256, 0, st>>>(t->log, t->dev, rc.chunk_base, rc.chunk_end, nn, rc.max_chunks,
                                                       rc.slots, rc.leader, rc.choice, rc.depth, rc.endpoint,
                                                       rc.matched, rc.n_dev, rc.idx_of_pos);
    B200R_LAUNCH_CHECK("k_commit");
    }
    return B200R_OK;
}

}  // namespace b200r

// =========================================================================
// C ABI
// =========================================================================
using namespace b200r;

extern "C" {

int b200r_abi_version(void) { return B200R_ABI_VERSION; }
const char *b200r_last_error(void) { return g_err; }
uint32_t b200r_pick(uint64_t seed, uint64_t seq, uint32_t n) { return pick(seed, seq, n); }

int b200r_set_claim_mode(int mode)
{
    if (mode < 0 || mode > 4) { set_error("b200r_set_claim_mode: 0 auto, 1 thread, 2 warp, 3 half warp, 4 quarter warp"); return B200R_E_INVALID; }
    g_claim_mode = mode;
    return B200R_OK;
}

int b200r_set_hash_mode(int mode)
{
    if (mode < 0 || mode > 1) { set_error("b200r_set_hash_mode: 0 streaming, 1 one tile per warp"); return B200R_E_INVALID; }
    g_hash_mode = mode;
    return B200R_OK;
}

int b200r_set_part_staged(int on)
{
    g_part_staged = on != 0;
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
                       d_hashes, max_chunks, (uint8_t *)d_scratch, p, nullptr, st, 0);
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
                       d_hashes, max_chunks, scratch, p, nullptr, st, 0);
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
    t->tab = nullptr; t->log = nullptr; t->wblk = nullptr; t->dev = nullptr;
    t->part_ready = false;
    for (int k = 0; k < PART_SETS; k++) t->part_has_finished[k] = false;
    t->one_host = t->one_host_dev = t->one_dev = nullptr;
    const size_t cap = (size_t)1 << cap_log2;
    cudaError_t e;
    if ((e = cudaMalloc(&t->tab, 2 * cap * sizeof(Head))) != cudaSuccess ||
        (e = cudaMalloc(&t->log, cap * sizeof(Rec))) != cudaSuccess ||
        (e = cudaMalloc(&t->wblk, (size_t)NWSLOT * sizeof(uint2))) != cudaSuccess ||
        (e = cudaMalloc(&t->dev, sizeof(TrieDev))) != cudaSuccess) {
        cudaFree(t->tab); cudaFree(t->log); cudaFree(t->wblk); cudaFree(t->dev);
        delete t;
        return cuda_fail(e, "cudaMalloc(trie)");
    }
    int rc = b200r_trie_clear(t, nullptr);
    if (rc) { b200r_trie_destroy(t); return rc; }
    B200R_CUDA(cudaStreamSynchronize(nullptr));
    *out = t;
    return B200R_OK;
}

int b200r_trie_destroy(b200r_trie *t)
{
    if (!t) return B200R_OK;
    if (t->part_ready) {
        cudaStream_t sts[] = { t->part_prep, t->part_send, t->part_route, t->part_ret };
        for (cudaStream_t s : sts) cudaStreamSynchronize(s);
        cudaEvent_t evs[] = { t->part_inputs, t->part_prepped, t->part_routed };
        for (cudaEvent_t e : evs) cudaEventDestroy(e);
        for (int k = 0; k < PART_SETS; k++) { cudaEventDestroy(t->part_sent[k]); cudaEventDestroy(t->part_finished[k]); }
        for (cudaStream_t s : sts) cudaStreamDestroy(s);
    }
    if (t->one_host) { cudaDeviceSynchronize(); cudaFreeHost(t->one_host); cudaFree(t->one_dev); }
    cudaFree(t->tab); cudaFree(t->log); cudaFree(t->wblk); cudaFree(t->dev);
    delete t;
    return B200R_OK;
}

int b200r_trie_clear(b200r_trie *t, void *stream)
{
    if (!t) { set_error("null trie"); return B200R_E_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    const size_t cap = (size_t)1 << t->cap_log2;
    // the head table starts empty, and so does the log: an all-zero record means "not written
    // yet" to a reader that got there ahead of the writer's store (trie.cuh)
    B200R_CUDA(cudaMemsetAsync(t->tab, 0, 2 * cap * sizeof(Head), st));
    B200R_CUDA(cudaMemsetAsync(t->log, 0, cap * sizeof(Rec), st));
    B200R_CUDA(cudaMemsetAsync(t->wblk, 0, (size_t)NWSLOT * sizeof(uint2), st));
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
        const int64_t cap = (int64_t)2 << t->cap_log2;
        int blocks = (int)((cap + 255) / 256);
        if (blocks > 148 * 16) blocks = 148 * 16;
        B200R_CUDA(cudaMemsetAsync(&t->dev->num_nodes, 0, 8, st));
        k_count_nodes<<<blocks, 256, 0, st>>>(t->tab, t->log, cap, t->dev);
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

int b200r_trie_status_async(b200r_trie *t, uint32_t *d_status_out, void *stream)
{
    if (!t || !d_status_out) { set_error("b200r_trie_status_async: null pointer"); return B200R_E_INVALID; }
    B200R_CUDA(cudaMemcpyAsync(d_status_out, &t->dev->status, 4, cudaMemcpyDeviceToDevice,
                               (cudaStream_t)stream));
    return B200R_OK;
}

int b200r_trie_retire_endpoints(b200r_trie *t, uint64_t bits, void *stream)
{
    if (!t) { set_error("null trie"); return B200R_E_INVALID; }
    if (!bits) return B200R_OK;
    g_launches += 1;
    k_clear_bits<<<148 * 8, 256, 0, (cudaStream_t)stream>>>(t->log, t->dev, (uint32_t)1 << t->cap_log2, bits);
    B200R_LAUNCH_CHECK("k_clear_bits");
    return B200R_OK;
}

int b200r_trie_set_epoch(b200r_trie *t, uint32_t epoch)
{
    if (!t) { set_error("null trie"); return B200R_E_INVALID; }
    t->epoch = epoch;
    return B200R_OK;
}

int b200r_trie_lookup_batch(b200r_trie *t, const uint64_t *d_hashes, const int32_t *d_chunk_base,
                            int64_t n, uint64_t avail_mask, int32_t *d_match_depth,
                            uint64_t *d_matched_mask, void *stream)
{
    if (!t || n < 0 || n > 0x7fffffff - 2048) { set_error("b200r_trie_lookup_batch: bad trie or n"); return B200R_E_INVALID; }
    if (n == 0) return B200R_OK;
    if (!d_hashes || !d_chunk_base || !d_match_depth || !d_matched_mask) { set_error("null pointer"); return B200R_E_INVALID; }
    WalkArgs a = WalkArgs();
    a.t = trie_ref(t);
    a.hashes = d_hashes; a.chunk_base = d_chunk_base; a.max_chunks = INT64_MAX;
    a.avail = avail_mask; a.depth_out = d_match_depth; a.matched_out = d_matched_mask;
    { ProfScope ps__(PK_LOOKUP, (cudaStream_t)stream); g_launches += 1;
    k_walk<W_LOOKUP, 32><<<walk_grid(n), WARPS_PER_CTA * 32, 0, (cudaStream_t)stream>>>(a, (int)n, nullptr);
    B200R_LAUNCH_CHECK("k_walk<lookup>");
    }
    return B200R_OK;
}

int b200r_trie_insert_batch(b200r_trie *t, const uint64_t *d_hashes, const int32_t *d_chunk_base,
                            int64_t n, const int32_t *d_endpoint, void *stream)
{
    if (!t || n < 0 || n > 0x7fffffff - 2048) { set_error("b200r_trie_insert_batch: bad trie or n"); return B200R_E_INVALID; }
    if (n == 0) return B200R_OK;
    if (!d_hashes || !d_chunk_base || !d_endpoint) { set_error("null pointer"); return B200R_E_INVALID; }
    WalkArgs a = WalkArgs();
    a.t = trie_ref(t);
    a.hashes = d_hashes; a.chunk_base = d_chunk_base; a.max_chunks = INT64_MAX;
    a.endpoint = d_endpoint;
    { ProfScope ps__(PK_INSERT, (cudaStream_t)stream); g_launches += 1;
    k_walk<W_INSERT, 32><<<walk_grid(n), WARPS_PER_CTA * 32, 0, (cudaStream_t)stream>>>(a, (int)n, nullptr);
    B200R_LAUNCH_CHECK("k_walk<insert>");
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
    const int phase = (flags & B200R_F_PREP_ONLY) ? 1 : ((flags & B200R_F_PREP_DONE) ? 2 : 0);
    return launch_hash(d_bytes, d_starts, d_lens, d_nchars, n, t->chunk_chars, flags,
                       (int32_t *)(sc + p.off_chunk_base), (uint64_t *)(sc + p.off_hashes),
                       max_chunks, sc, p, t->dev, (cudaStream_t)stream, phase);
}

static int route_from(b200r_trie *t, const uint64_t *hashes, const int32_t *chunk_base, int64_t n,
                      int64_t max_chunks, uint64_t avail_mask, uint64_t seed, uint64_t seq0,
                      int32_t *d_endpoint, int32_t *d_match_depth, uint64_t *d_matched_mask,
                      uint8_t *sc, const Plan &p, cudaStream_t st)
{
    RouteCore rc = RouteCore();
    rc.hashes = hashes; rc.chunk_base = chunk_base;
    rc.n = n; rc.max_chunks = max_chunks;
    rc.avail = avail_mask; rc.seed = seed; rc.seq0 = seq0;
    rc.slots = (uint32_t *)(sc + p.off_slots);
    rc.leader = (int32_t *)(sc + p.off_leader);
    rc.choice = (int32_t *)(sc + p.off_choice);
    rc.dpre = (int32_t *)(sc + p.off_dpre);
    rc.cand = (uint64_t *)(sc + p.off_cand);
    rc.res = (unsigned long long *)(sc + p.off_res);
    rc.endpoint = d_endpoint; rc.depth = d_match_depth; rc.matched = d_matched_mask;
    // a thread per request while prompts are short on average, a warp per request beyond
    rc.claim_kind = claim_kind_for(n, max_chunks);
    return launch_route_core(t, rc, st);
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
    uint8_t *sc = (uint8_t *)d_scratch;
    return route_from(t, (const uint64_t *)(sc + p.off_hashes), (const int32_t *)(sc + p.off_chunk_base),
                      n, max_chunks, avail_mask, seed, seq0, d_endpoint, d_match_depth, d_matched_mask,
                      sc, p, (cudaStream_t)stream);
}

int b200r_prefix_route_hashes(b200r_trie *t, const uint64_t *d_hashes, const int32_t *d_chunk_base,
                              int64_t n, int64_t max_chunks, uint64_t avail_mask, uint64_t seed,
                              uint64_t seq0, int32_t *d_endpoint, int32_t *d_match_depth,
                              uint64_t *d_matched_mask, void *d_scratch, size_t scratch_bytes,
                              void *stream)
{
    if (!t || n < 0 || max_chunks < 0) { set_error("b200r_prefix_route_hashes: bad argument"); return B200R_E_INVALID; }
    if (n == 0) return B200R_OK;
    if (!d_hashes || !d_chunk_base || !d_endpoint || !d_match_depth || !d_matched_mask) { set_error("null pointer"); return B200R_E_INVALID; }
    const Plan p = make_plan(n, max_chunks, B200R_F_ALL_ASCII, true);
    if (!d_scratch || scratch_bytes < p.total) {
        set_error("scratch too small: need %zu bytes, got %zu", p.total, scratch_bytes);
        return B200R_E_SCRATCH;
    }
    return route_from(t, d_hashes, d_chunk_base, n, max_chunks, avail_mask, seed, seq0, d_endpoint,
                      d_match_depth, d_matched_mask, (uint8_t *)d_scratch, p, (cudaStream_t)stream);
}

// a5 for ONE request with host buffers in and out (see k_route_one).  Synchronous: one launch
// and one stream synchronisation.  Returns 0 = routed, 1 = not taken (prompt not ASCII, longer
// than the path's staging buffer, or chunk size not a multiple of 16): route it as a batch of
// one.  The staging buffer (pinned, host-mapped) and 200 KB of device scratch are allocated on
// the table's first call.
int b200r_prefix_route_one(b200r_trie *t, const uint8_t *h_prompt, int64_t len_bytes,
                           uint64_t avail_mask, uint64_t seed, uint64_t seq, int32_t *endpoint,
                           int32_t *match_depth, uint64_t *matched_mask, uint32_t *status,
                           void *stream)
{
    if (!t || len_bytes < 0 || (len_bytes && !h_prompt) || !endpoint) { set_error("b200r_prefix_route_one: bad argument"); return B200R_E_INVALID; }
    const int cc = t->chunk_chars;
    if ((cc & 15) || cc > ONE_PASS || len_bytes > (int64_t)ONE_MAX_CHUNKS * cc) return 1;
    cudaStream_t st = (cudaStream_t)stream;
    if (!t->one_host) {
        const size_t hb = (size_t)ONE_MAX_CHUNKS * cc + 256;
        B200R_CUDA(cudaHostAlloc((void **)&t->one_host, hb, cudaHostAllocMapped));
        B200R_CUDA(cudaHostGetDevicePointer((void **)&t->one_host_dev, t->one_host, 0));
        const size_t db = (size_t)ONE_MAX_CHUNKS * 12 + 256;
        cudaError_t e = cudaMalloc((void **)&t->one_dev, db);
        if (e != cudaSuccess) { cudaFreeHost(t->one_host); t->one_host = nullptr; set_error("b200r_prefix_route_one: %s", cudaGetErrorString(e)); return B200R_E_NOMEM; }
    }
    // layout: host = [result block 64 B | prompt]; device = [hashes | slots | chunk_base, dpre, cand]
    OneResult *res = (OneResult *)t->one_host;
    uint8_t *stage = t->one_host + 64;
    memcpy(stage, h_prompt, (size_t)len_bytes);
    memset(stage + len_bytes, 0, 16);
    res->status = 0;
    unsigned long long tag;
    int rc = next_epoch(t, st, &tag);
    if (rc) return rc;
    uint64_t *hashes = (uint64_t *)t->one_dev;
    uint32_t *slots = (uint32_t *)(t->one_dev + (size_t)ONE_MAX_CHUNKS * 8);
    uint8_t *tail = t->one_dev + (size_t)ONE_MAX_CHUNKS * 12;
    WalkArgs a = WalkArgs();
    a.t = trie_ref(t);
    a.hashes = hashes; a.chunk_base = (int32_t *)tail; a.max_chunks = ONE_MAX_CHUNKS;
    a.avail = avail_mask; a.tag = tag; a.slots = slots;
    a.dpre = (int32_t *)(tail + 8); a.cand = (uint64_t *)(tail + 16);
    g_launches += 1;
    k_route_one<<<1, 256, 0, st>>>(a, t->dev, t->one_host_dev + 64, (int)len_bytes, cc, seed, seq,
                                   (int32_t *)tail, hashes, (OneResult *)t->one_host_dev);
    B200R_LAUNCH_CHECK("k_route_one");
    B200R_CUDA(cudaStreamSynchronize(st));
    const uint32_t s = *(volatile uint32_t *)&res->status;
    if (!(s & ONE_DONE)) { set_error("b200r_prefix_route_one: the kernel left no result"); return B200R_E_CUDA; }
    if (s & ONE_NONASCII) return 1;
    *endpoint = res->endpoint;
    if (match_depth) *match_depth = res->depth;
    if (matched_mask) *matched_mask = res->matched;
    if (status) *status = s & 0xFFFFu;
    return B200R_OK;
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

// ---------------------------------------------------------------------------
// The three steps of a routed batch (prep | hash | route) pipelined over consecutive batches
// natively: stream and event handling in Python costs more host time per batch than the GPU
// needs for the batch.
constexpr int64_t PIPE_SERIAL_MAX = 4096;   // batches up to this size skip the stream pipeline

struct b200r_pipe {
    b200r_trie *t;
    int depth;
    long long calls;
    cudaStream_t prep, hash, route;
    cudaEvent_t prepped[8], hashed[8], done[8], inputs;
    bool has_done[8];
};

int b200r_pipe_create(b200r_trie *t, int32_t depth, b200r_pipe **out)
{
    if (!t || !out || depth < 1 || depth > 8) { set_error("b200r_pipe_create: trie, 1 <= depth <= 8"); return B200R_E_INVALID; }
    b200r_pipe *p = new b200r_pipe();
    p->t = t; p->depth = depth; p->calls = 0;
    // The route step (small, latency-bound kernels) and the prep step get the highest stream
    // priority, the hashing kernel (thousands of CTAs that fill every SM) the lowest: without
    // it the block scheduler hands out the hashing kernel's CTAs first and the route step of the
    // batch before only starts when they are nearly through -- the steps would not overlap.
    int least = 0, greatest = 0;
    B200R_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    // route steps first; prep and hashing of the following batches fill in behind them
    B200R_CUDA(cudaStreamCreateWithPriority(&p->prep, cudaStreamNonBlocking, least));
    B200R_CUDA(cudaStreamCreateWithPriority(&p->hash, cudaStreamNonBlocking, least));
    B200R_CUDA(cudaStreamCreateWithPriority(&p->route, cudaStreamNonBlocking, greatest));
    B200R_CUDA(cudaEventCreateWithFlags(&p->inputs, cudaEventDisableTiming));
    for (int k = 0; k < depth; k++) {
        B200R_CUDA(cudaEventCreateWithFlags(&p->prepped[k], cudaEventDisableTiming));
        B200R_CUDA(cudaEventCreateWithFlags(&p->hashed[k], cudaEventDisableTiming));
        B200R_CUDA(cudaEventCreateWithFlags(&p->done[k], cudaEventDisableTiming));
        p->has_done[k] = false;
    }
    *out = p;
    return B200R_OK;
}

int b200r_pipe_destroy(b200r_pipe *p)
{
    if (!p) return B200R_OK;
    cudaStreamSynchronize(p->prep); cudaStreamSynchronize(p->hash); cudaStreamSynchronize(p->route);
    for (int k = 0; k < p->depth; k++) {
        cudaEventDestroy(p->prepped[k]); cudaEventDestroy(p->hashed[k]); cudaEventDestroy(p->done[k]);
    }
    cudaEventDestroy(p->inputs);
    cudaStreamDestroy(p->prep); cudaStreamDestroy(p->hash); cudaStreamDestroy(p->route);
    delete p;
    return B200R_OK;
}

void *b200r_pipe_prep_stream(b200r_pipe *p) { return p ? (void *)p->prep : nullptr; }

int b200r_pipe_copy_h2d(b200r_pipe *p, void *d_dst, const void *h_src, size_t bytes)
{
    if (!p || (bytes && (!d_dst || !h_src))) { set_error("b200r_pipe_copy_h2d: null pointer"); return B200R_E_INVALID; }
    if (bytes) B200R_CUDA(cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, p->prep));
    return B200R_OK;
}

int b200r_pipe_begin(b200r_pipe *p)
{
    if (!p) { set_error("null pipeline"); return B200R_E_INVALID; }
    const int k = (int)(p->calls % p->depth);
    if (p->has_done[k]) {
        B200R_CUDA(cudaStreamWaitEvent(p->prep, p->done[k], 0));
        p->has_done[k] = false;               // waited for: b200r_pipe_route need not wait again
    }
    return B200R_OK;
}

int b200r_pipe_route(b200r_pipe *p, const uint8_t *d_bytes, const int64_t *d_starts,
                     const int32_t *d_lens, const int32_t *d_nchars, int64_t n, int64_t max_chunks,
                     uint32_t flags, uint64_t avail_mask, uint64_t seed, uint64_t seq0,
                     int32_t *d_endpoint, int32_t *d_match_depth, uint64_t *d_matched_mask,
                     void *d_scratch, size_t scratch_bytes, int32_t inputs_on_prep_stream, void *stream)
{
    if (!p) { set_error("null pipeline"); return B200R_E_INVALID; }
    b200r_trie *t = p->t;
    int rc = check_prompt_args(d_bytes, d_starts, d_lens, n, max_chunks, t->chunk_chars);
    if (rc) return rc;
    if (n == 0) return B200R_OK;
    if (!d_endpoint || !d_match_depth || !d_matched_mask) { set_error("null output pointer"); return B200R_E_INVALID; }
    const Plan pl = make_plan(n, max_chunks, flags, true);
    if (!d_scratch || scratch_bytes < pl.total) {
        set_error("scratch too small: need %zu bytes, got %zu", pl.total, scratch_bytes);
        return B200R_E_SCRATCH;
    }
    cudaStream_t cur = (cudaStream_t)stream;
    const int k = (int)(p->calls % p->depth);
    p->calls++;
    uint8_t *sc = (uint8_t *)d_scratch;
    int32_t *chunk_base = (int32_t *)(sc + pl.off_chunk_base);
    uint64_t *hashes = (uint64_t *)(sc + pl.off_hashes);
    if (n <= PIPE_SERIAL_MAX) {
        // Small batch: the GPU needs less time for it than the host needs to hand it to three
        // streams (a dozen event calls), so all six kernels go to the route stream in order.
        // Scratch k: every earlier user's stages precede its route step, which is on this stream.
        if (inputs_on_prep_stream) {                   // the copies were put on the prep stream
            B200R_CUDA(cudaEventRecord(p->prepped[k], p->prep));
            B200R_CUDA(cudaStreamWaitEvent(p->route, p->prepped[k], 0));
        }
        B200R_CUDA(cudaEventRecord(p->inputs, cur));   // inputs produced / outputs last used there
        B200R_CUDA(cudaStreamWaitEvent(p->route, p->inputs, 0));
        rc = launch_hash(d_bytes, d_starts, d_lens, d_nchars, n, t->chunk_chars, flags, chunk_base, hashes,
                         max_chunks, sc, pl, t->dev, p->route, 0);
        if (rc) return rc;
        rc = route_from(t, hashes, chunk_base, n, max_chunks, avail_mask, seed, seq0, d_endpoint,
                        d_match_depth, d_matched_mask, sc, pl, p->route);
        if (rc) return rc;
        B200R_CUDA(cudaEventRecord(p->done[k], p->route));
        B200R_CUDA(cudaStreamWaitEvent(cur, p->done[k], 0));
        p->has_done[k] = true;
        return B200R_OK;
    }
    if (p->has_done[k]) B200R_CUDA(cudaStreamWaitEvent(p->prep, p->done[k], 0));   // scratch k is free again
    if (!inputs_on_prep_stream) {
        B200R_CUDA(cudaEventRecord(p->inputs, cur));
        B200R_CUDA(cudaStreamWaitEvent(p->prep, p->inputs, 0));
    }
    rc = launch_hash(d_bytes, d_starts, d_lens, d_nchars, n, t->chunk_chars, flags, chunk_base, hashes,
                     max_chunks, sc, pl, t->dev, p->prep, 1);
    if (rc) return rc;
    B200R_CUDA(cudaEventRecord(p->prepped[k], p->prep));
    B200R_CUDA(cudaStreamWaitEvent(p->hash, p->prepped[k], 0));
    rc = launch_hash(d_bytes, d_starts, d_lens, d_nchars, n, t->chunk_chars, flags, chunk_base, hashes,
                     max_chunks, sc, pl, t->dev, p->hash, 2);
    if (rc) return rc;
    B200R_CUDA(cudaEventRecord(p->hashed[k], p->hash));
    // the route step runs on the pipeline's own high-priority stream, in call order; the
    // outputs may have been in use on the caller's stream, and the caller's stream sees them
    // when the step is done
    B200R_CUDA(cudaStreamWaitEvent(p->route, p->hashed[k], 0));
    if (!inputs_on_prep_stream) B200R_CUDA(cudaStreamWaitEvent(p->route, p->inputs, 0));
    else {
        B200R_CUDA(cudaEventRecord(p->inputs, cur));
        B200R_CUDA(cudaStreamWaitEvent(p->route, p->inputs, 0));
    }
    rc = route_from(t, hashes, chunk_base, n, max_chunks, avail_mask, seed, seq0, d_endpoint,
                    d_match_depth, d_matched_mask, sc, pl, p->route);
    if (rc) return rc;
    B200R_CUDA(cudaEventRecord(p->done[k], p->route));
    B200R_CUDA(cudaStreamWaitEvent(cur, p->done[k], 0));
    p->has_done[k] = true;
    return B200R_OK;
}

int b200r_trie_dump(b200r_trie *t, uint64_t *d_out, int64_t max_nodes, int64_t *count,
                    uint64_t *root_mask, void *stream)
{
    if (!t || max_nodes < 0 || (max_nodes > 0 && !d_out)) { set_error("b200r_trie_dump: bad argument"); return B200R_E_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    B200R_CUDA(cudaMemsetAsync(&t->dev->dump_count, 0, 8, st));
    const int64_t cap = (int64_t)2 << t->cap_log2;
    int blocks = (int)((cap + 255) / 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_dump<<<blocks, 256, 0, st>>>(t->tab, t->log, cap, t->dev, d_out, max_nodes);
    B200R_LAUNCH_CHECK("k_dump");
    TrieDev h;
    B200R_CUDA(cudaMemcpyAsync(&h, t->dev, sizeof h, cudaMemcpyDeviceToHost, st));
    B200R_CUDA(cudaStreamSynchronize(st));
    if (count) *count = (int64_t)h.dump_count;
    if (root_mask) *root_mask = h.root_mask;
    return B200R_OK;
}

int b200r_trie_restore(b200r_trie *t, const uint64_t *d_entries, int64_t count,
                       uint64_t root_mask, void *stream)
{
    if (!t || count < 0 || (count > 0 && !d_entries)) { set_error("b200r_trie_restore: bad argument"); return B200R_E_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    TrieDev h;
    B200R_CUDA(cudaMemcpyAsync(&h, t->dev, sizeof h, cudaMemcpyDeviceToHost, st));
    B200R_CUDA(cudaStreamSynchronize(st));
    if (h.log_count != 0) {
        set_error("b200r_trie_restore: the table is not empty (node identity is the position in "
                  "the table's log, so a dump can only be restored into an empty table)");
        return B200R_E_INVALID;
    }
    const int64_t work = count > 0 ? count : 1;
    g_launches += 1;
    k_restore<<<(int)((work + 255) / 256), 256, 0, st>>>(trie_ref(t), d_entries, count, root_mask);
    B200R_LAUNCH_CHECK("k_restore");
    return B200R_OK;
}

}  // extern "C"
