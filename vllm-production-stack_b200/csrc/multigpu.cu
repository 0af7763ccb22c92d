// multigpu.cu -- prefix-aware routing of one global batch over several GPUs with a
// PARTITIONED table (one process per GPU; no counterpart in the reference, whose router
// replicas share nothing).  C ABI: b200r_part_* (include/b200_router.h).  The replicated-table
// variant needs no kernels of its own: it all-gathers chunk hashes on the host side (NCCL) and
// routes the global batch with b200r_prefix_route_hashes (prefix.cu).
#include <algorithm>

#include "internal.cuh"

namespace b200r {

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


// =========================================================================
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
    unsigned long long in_req[PART_SETS][P2P_MAX_WORLD], in_chunk[PART_SETS][P2P_MAX_WORLD],
        in_n[PART_SETS][P2P_MAX_WORLD], res[PART_SETS][P2P_MAX_WORLD];
};
static_assert(sizeof(PartFlags) <= 8192, "flag block");
struct PartRes { int32_t ep, depth; uint64_t matched; };
struct PartDyn {
    P2PDyn base;                              // seq_base, total_n (my_off stays 0)
    int32_t n_owned, c_owned;
    uint32_t reqbase[P2P_MAX_WORLD + 1], chunkbase[P2P_MAX_WORLD + 1], off[P2P_MAX_WORLD + 1];
    unsigned int ticket_bulk, ticket_res;     // "last CTA publishes" counters (self-resetting)
};

// Every CTA calls this after its last peer store: one system-scope fence per CTA, then a
// ticket; the CTA that draws the last ticket knows that all peer stores of the grid are
// visible system-wide and may publish the arrival flags (returns true for it, block-wide).
__device__ __forceinline__ bool last_cta_after_peer_stores(unsigned int *ticket)
{
    __shared__ int s_last;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        const unsigned int t = atomicAdd(ticket, 1u);
        s_last = (t == gridDim.x - 1);
        if (s_last) *ticket = 0;              // ready for the next launch
    }
    __syncthreads();
    return s_last != 0;
}

__device__ __forceinline__ int part_owner(const uint64_t *hashes, int b, int C, int world)
{
    if (C == 0) return 0;                     // empty prompt: touches only a root
    return part_owner_of(hashes[b], world);
}

constexpr int PART_BLOCK = 256;

// Where does every request (and its run of chunk hashes) go in its owner's buffer?  Requests
// need not arrive in order: the owner identifies a request by its GLOBAL position (source
// offset + local index), which is what "first toucher", leader links and the tie-break counter
// are expressed in.  So a slot in the (source, owner) segment is simply taken with one packed
// atomicAdd per warp and owner {requests << 32 | chunks} -- no counting pass, no scan.  The
// {local index, chunk offset in the segment} words and the hashes are first compacted into
// LOCAL staging segments (one per owner, fixed capacity) and then shipped by k_part_bulk with
// full-line 16-byte vector stores: NVLink moves the same bytes several times faster in
// 128-byte lines than as 8-byte words or ragged ~170-byte runs (measured at 2 GPUs: 47 us per
// shard for direct scattered peer stores).
__global__ void __launch_bounds__(PART_BLOCK)
k_part_place(const int32_t *__restrict__ chunk_base, const uint64_t *__restrict__ hashes, int n,
             int world, unsigned long long *__restrict__ totals, int64_t cap, int64_t rcap,
             unsigned long long *__restrict__ info_stage, uint32_t *__restrict__ dest_slot,
             uint64_t *__restrict__ hash_stage)
{
    __shared__ int s_ex[PART_BLOCK / 32][32];
    __shared__ unsigned long long s_dst[PART_BLOCK];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int i = blockIdx.x * PART_BLOCK + threadIdx.x;
    int b = 0, C = 0, o = -1;
    if (i < n) {
        b = chunk_base[i];
        C = chunk_base[i + 1] - b;
        o = part_owner(hashes, b, C, world);
    }
    uint32_t rreq = 0, rchunk = 0;
    for (int d = 0; d < world; d++) {
        const unsigned m = __ballot_sync(0xffffffffu, o == d);
        if (!m) continue;                                   // warp-uniform
        int v = (o == d) ? C : 0, inc = v;
        for (int sft = 1; sft < 32; sft <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, sft);
            if (lane >= sft) inc += t;
        }
        const int tot = __shfl_sync(0xffffffffu, inc, 31);
        unsigned long long base = 0;
        if (lane == 0)
            base = atomicAdd(&totals[d], ((unsigned long long)__popc(m) << 32) | (unsigned long long)(uint32_t)tot);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (o == d) {
            rreq = (uint32_t)(base >> 32) + (uint32_t)__popc(m & ((1u << lane) - 1));
            rchunk = (uint32_t)base + (uint32_t)(inc - v);
        }
    }
    // chunk hashes -> local staging.  The 32 requests of a warp are consecutive in the batch, so
    // their hashes are ONE contiguous source range: the warp copies it flat (every lane busy,
    // four loads in flight) and looks the destination up by request (search over 32 prefixes).
    int ex = C;
    for (int sft = 1; sft < 32; sft <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, ex, sft);
        if (lane >= sft) ex += t;
    }
    const int T = __shfl_sync(0xffffffffu, ex, 31);
    ex -= C;
    s_ex[w][lane] = ex;
    s_dst[threadIdx.x] = 0;
    if (o >= 0) {
        dest_slot[i] = ((uint32_t)o << 24) | rreq;
        s_dst[threadIdx.x] = (unsigned long long)o * (unsigned long long)cap + rchunk - (unsigned)ex;
        info_stage[(size_t)o * rcap + rreq] =
            (unsigned long long)(uint32_t)i | ((unsigned long long)rchunk << 32);
    }
    const int cb0 = __shfl_sync(0xffffffffu, b, 0);          // lane 0's request exists if any does
    __syncwarp();
    const uint64_t *src = hashes + cb0;
    const int *pex = s_ex[w];
    const unsigned long long *pdst = s_dst + w * 32;
    for (int j0 = 0; j0 < T; j0 += 128) {
        uint64_t v[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int j = j0 + u * 32 + lane;
            v[u] = j < T ? __ldcs(src + j) : 0;
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int j = j0 + u * 32 + lane;
            if (j >= T) continue;
            int r = 0;                                       // largest r with ex[r] <= j
#pragma unroll
            for (int sft = 16; sft; sft >>= 1)
                if (pex[r + sft] <= j) r += sft;
            hash_stage[pdst[r] + (unsigned)j] = v[u];
        }
    }
}

// grid-strided copy of 16-byte vectors, four loads in flight per thread
__device__ __forceinline__ void copy_vec4(uint4 *__restrict__ dst, const uint4 *__restrict__ src,
                                          int64_t cnt, int64_t tid, int64_t nthr)
{
    for (int64_t x0 = tid; x0 < cnt; x0 += 4 * nthr) {
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; u++)
            if (x0 + u * nthr < cnt) v[u] = __ldcs(src + x0 + u * nthr);
#pragma unroll
        for (int u = 0; u < 4; u++)
            if (x0 + u * nthr < cnt) dst[x0 + u * nthr] = v[u];
    }
}

// staging segments -> the owners' buffers: 16-byte vectors, whole lines.  Persistent grid, one
// system-scope fence per CTA at the end (MEMBAR.SYS costs microseconds: not once per wave);
// the last CTA to finish publishes the arrival flags (no separate launch).
__global__ void __launch_bounds__(256)
k_part_bulk(const uint64_t *__restrict__ hash_stage, const unsigned long long *__restrict__ info_stage,
            unsigned long long *__restrict__ totals, int world, int my_rank,
            const unsigned long long *__restrict__ peer_bases, size_t hash_off, size_t info_off,
            int64_t cap, int64_t rcap, unsigned int *ticket, int set, unsigned long long epoch,
            unsigned int n)
{
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t nthr = (int64_t)gridDim.x * blockDim.x;
    for (int k = 0; k < world; k++) {
        const int o = (my_rank + 1 + k) % world;       // start with the next rank: spread the links
        const unsigned long long tt = totals[o];
        const int nreq = (int)(tt >> 32), nchunk = (int)(uint32_t)tt;
        const uint4 *hs = (const uint4 *)(hash_stage + (size_t)o * cap);
        uint4 *hd = (uint4 *)((uint64_t *)(uintptr_t)(peer_bases[o] + hash_off) + (size_t)my_rank * cap);
        copy_vec4(hd, hs, (nchunk + 1) / 2, tid, nthr);
        const uint4 *is = (const uint4 *)(info_stage + (size_t)o * rcap);
        uint4 *id = (uint4 *)((unsigned long long *)(uintptr_t)(peer_bases[o] + info_off) + (size_t)my_rank * rcap);
        copy_vec4(id, is, (nreq + 1) / 2, tid, nthr);
    }
    if (last_cta_after_peer_stores(ticket) && (int)threadIdx.x < world) {
        const int d = threadIdx.x;
        const unsigned long long tt = totals[d];
        PartFlags *f = (PartFlags *)(uintptr_t)peer_bases[d];
        *(volatile unsigned long long *)&f->in_n[set][my_rank] = (epoch << 32) | n;
        *(volatile unsigned long long *)&f->in_chunk[set][my_rank] = (epoch << 32) | (uint32_t)tt;
        *(volatile unsigned long long *)&f->in_req[set][my_rank] = (epoch << 32) | (uint32_t)(tt >> 32);
        __threadfence_system();
        totals[d] = 0;                                 // every CTA has read it: ready for the next batch
    }
}

// Arrival flags for every owner.  Runs behind the kernels that stored into the owners' buffers
// (stream order: their stores are complete when this starts).  `totals` = the packed
// {requests << 32 | chunks} per owner that the placement left (NULL: an empty batch); reset
// here for the batch after next.
__global__ void k_part_publish_in(const unsigned long long *__restrict__ peer_bases, int world,
                                  int my_rank, int set, unsigned long long epoch, unsigned int n,
                                  unsigned long long *__restrict__ totals)
{
    const int d = threadIdx.x;
    if (d >= world) return;
    const unsigned long long tt = totals ? totals[d] : 0ULL;
    PartFlags *f = (PartFlags *)(uintptr_t)peer_bases[d];
    __threadfence_system();
    *(volatile unsigned long long *)&f->in_n[set][my_rank] = (epoch << 32) | n;
    *(volatile unsigned long long *)&f->in_chunk[set][my_rank] = (epoch << 32) | (uint32_t)tt;
    *(volatile unsigned long long *)&f->in_req[set][my_rank] = (epoch << 32) | (uint32_t)(tt >> 32);
    __threadfence_system();
    if (totals) totals[d] = 0;
}

// Wait for all sources.  ONE small CTA spins on the arrival flags (local memory): a grid of
// spinning CTAs would hold SM slots that the next batch's send phase needs while the slowest
// peer is still on its way.  Leaves the batch's bookkeeping in `dyn`.
__global__ void k_part_wait_meta(const PartFlags *__restrict__ mine, int set, unsigned long long epoch,
                                 int world, PartDyn *dyn, const PartDyn *prev, TrieDev *dev)
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
        // global sequence number of this batch's request 0: the previous batch's, plus its size
        dyn->base.seq_base = prev->base.seq_base + prev->base.total_n;
    }
    // no fence: what the sources wrote before their flags is read by LATER launches only
}

// The requests this rank owns, in rank order of their sources: where each one's chunk hashes
// lie in the inbound buffer (they are routed from there; no copy), its global position, and the
// inverse map.  A request's run ends where the next slot of its (source, owner) segment begins
// (slots and chunk offsets were handed out together, so both are monotone within a segment).
__global__ void __launch_bounds__(256)
k_part_gather(const uint8_t *__restrict__ my_base, size_t info_off, int64_t cap, int64_t rcap,
              int world, const PartDyn *__restrict__ dyn, int32_t *__restrict__ chunk_base,
              int32_t *__restrict__ chunk_end, uint32_t *__restrict__ pos_owned,
              uint32_t *__restrict__ idx_of_pos)
{
    __shared__ uint32_t s_req[P2P_MAX_WORLD + 1], s_chunk[P2P_MAX_WORLD + 1], s_n[P2P_MAX_WORLD + 1];
    for (int k = threadIdx.x; k <= world; k += blockDim.x) {
        s_req[k] = dyn->reqbase[k]; s_chunk[k] = dyn->chunkbase[k]; s_n[k] = dyn->off[k];
    }
    __syncthreads();
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t step = (int64_t)gridDim.x * blockDim.x;
    const unsigned long long *info = (const unsigned long long *)(my_base + info_off);
    const int64_t n_owned = s_req[world];
    for (int64_t j = tid; j < n_owned; j += step) {
        int k = 0;
        while (k + 1 < world && j >= s_req[k + 1]) k++;
        const uint32_t r = (uint32_t)(j - s_req[k]), nreq = s_req[k + 1] - s_req[k];
        const unsigned long long v = info[(size_t)k * rcap + r];
        const uint32_t seg = (uint32_t)(v >> 32);
        const uint32_t nxt = (r + 1 < nreq) ? (uint32_t)(info[(size_t)k * rcap + r + 1] >> 32)
                                            : s_chunk[k + 1] - s_chunk[k];
        chunk_base[j] = (int32_t)((uint32_t)k * (uint32_t)cap + seg);
        chunk_end[j] = (int32_t)((uint32_t)k * (uint32_t)cap + nxt);
        const uint32_t gp = s_n[k] + (uint32_t)v;       // position in the global (rank-major) batch
        pos_owned[j] = gp;
        idx_of_pos[gp] = (uint32_t)j;
    }
}

__global__ void __launch_bounds__(256)
k_part_push_results(PartDyn *dyn, int world, int my_rank,
                    const int32_t *__restrict__ ep, const int32_t *__restrict__ depth,
                    const uint64_t *__restrict__ matched,
                    const unsigned long long *__restrict__ peer_bases,
                    size_t res_off, int64_t rcap, int set, unsigned long long epoch)
{
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < dyn->n_owned;
         j += (int64_t)gridDim.x * blockDim.x) {
        int q = 0;
        while (q + 1 < world && j >= dyn->reqbase[q + 1]) q++;
        PartRes r;
        r.ep = ep[j]; r.depth = depth[j]; r.matched = matched[j];
        ((PartRes *)(uintptr_t)(peer_bases[q] + res_off))[(size_t)my_rank * rcap + (j - dyn->reqbase[q])] = r;
    }
    if (last_cta_after_peer_stores(&dyn->ticket_res) && (int)threadIdx.x < world) {
        const int q = threadIdx.x;
        PartFlags *f = (PartFlags *)(uintptr_t)peer_bases[q];
        *(volatile unsigned long long *)&f->res[set][my_rank] =
            (epoch << 32) | (dyn->reqbase[q + 1] - dyn->reqbase[q]);
        __threadfence_system();
    }
}

// wait for all owners' results (one small CTA, see k_part_wait_meta)
__global__ void k_part_wait_res(const PartFlags *__restrict__ mine, int set, unsigned long long epoch,
                                int world, TrieDev *dev)
{
    const int q = threadIdx.x;
    if (q < world) {
        unsigned long long v;
        if (!p2p_spin(&mine->res[set][q], epoch, &v)) atomicOr(&dev->status, B200R_S_PEER_TIMEOUT);
    }
}

// results back to request order
__global__ void __launch_bounds__(256)
k_part_scatter(const uint32_t *__restrict__ dest_slot, const uint8_t *__restrict__ my_base,
               size_t res_off, int64_t rcap, int n, int32_t *__restrict__ endpoint,
               int32_t *__restrict__ depth, uint64_t *__restrict__ matched)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t ds = dest_slot[i];
    const PartRes r = ((const PartRes *)(my_base + res_off))[(size_t)(ds >> 24) * rcap + (ds & 0xFFFFFFu)];
    endpoint[i] = r.ep; depth[i] = r.depth; matched[i] = r.matched;
}

}  // namespace b200r

using namespace b200r;

extern "C" {

// ------------------------------------------------------ multi-GPU, partitioned
struct PartLayout {
    size_t symm_total, hash_off, info_off, res_off, hash_set, info_set, res_set;
    // local work buffer
    size_t work_total, w_dyn, w_totals, w_out_ptr, w_dest_slot, w_hash_stage, w_info_stage, w_chunk_end, w_pos_owned, w_idx_of_pos,
        w_ep, w_depth, w_matched, w_plan;
    int64_t n_max, c_max;
    int nblk_max;
};

static PartLayout part_layout(int world, int64_t cap, int64_t rcap)
{
    PartLayout L;
    L.hash_off = 8192;                                  // the flag block comes first
    L.hash_set = (size_t)world * cap * 8;
    L.info_off = L.hash_off + PART_SETS * L.hash_set;
    L.info_set = (size_t)world * rcap * 8;
    L.res_off = L.info_off + PART_SETS * L.info_set;
    L.res_set = (size_t)world * rcap * sizeof(PartRes);
    L.symm_total = align_up(L.res_off + PART_SETS * L.res_set);
    L.n_max = (int64_t)world * rcap;
    L.c_max = (int64_t)world * cap;
    L.nblk_max = (int)((rcap + PART_BLOCK - 1) / PART_BLOCK);
    size_t o = 0;
    L.w_dyn = o;       o += PART_SETS * align_up(sizeof(PartDyn));          // one per buffer set
    L.w_totals = o;    o += PART_SETS * align_up((size_t)world * 8);               // one per buffer set
    L.w_out_ptr = o;   o += PART_SETS * align_up((size_t)rcap * 8);
    L.w_dest_slot = o; o += PART_SETS * align_up((size_t)rcap * 4);     // one per buffer set
    L.w_hash_stage = o; o += align_up((size_t)world * (cap + 2) * 8);   // one fixed segment per owner
    L.w_info_stage = o; o += align_up((size_t)world * (rcap + 2) * 8);
    L.w_chunk_end = o; o += align_up((size_t)L.n_max * 4);
    L.w_pos_owned = o; o += align_up((size_t)L.n_max * 4);
    L.w_idx_of_pos = o; o += align_up((size_t)L.n_max * 4);
    L.w_ep = o;        o += PART_SETS * align_up((size_t)L.n_max * 4);      // results: one per buffer set
    L.w_depth = o;     o += PART_SETS * align_up((size_t)L.n_max * 4);
    L.w_matched = o;   o += PART_SETS * align_up((size_t)L.n_max * 8);
    L.w_plan = o;      o += make_plan(L.n_max, L.c_max, B200R_F_ALL_ASCII, true).total;
    L.work_total = o + 256;
    return L;
}

static inline PartDyn *part_dyn(uint8_t *wk, const PartLayout &L, int set)
{
    return (PartDyn *)(wk + L.w_dyn + (size_t)set * align_up(sizeof(PartDyn)));
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

// Send phase.  With a prep stream, the per-request counting + scan of the hash stage runs there
// (it is a chain of small kernels; next to the previous batch's hashing kernel instead of after
// it) and `st` picks up at the hashing kernel.
//   ASCII batches (the rule): the prep kernel places the requests (PartPlace) and the hashing
//   kernel stores every chunk hash straight into the owner's inbound buffer over NVLink peer
//   memory -- the hashing is fused with the dispatch; a one-CTA kernel behind it publishes the
//   arrival flags.  No copy of the hashes is made on the sending side.
//   Other batches: hashes in CSR order, then k_part_place (compaction into per-owner staging
//   segments) and k_part_bulk (shipment).
static int part_send_impl(b200r_trie *t, const uint8_t *d_bytes, const int64_t *d_starts,
                          const int32_t *d_lens, const int32_t *d_nchars, int64_t n, int64_t max_chunks,
                          uint32_t flags, const uint64_t *d_peer_bases, int32_t world, int32_t rank,
                          int64_t cap, int64_t rcap, uint32_t epoch, void *d_work, void *d_scratch,
                          size_t scratch_bytes, cudaStream_t prep_st, cudaEvent_t prepped, cudaStream_t st)
{
    int rc = part_check(world, rank, cap, rcap, d_peer_bases, d_work);
    if (rc) return rc;
    if (!t) { set_error("null trie"); return B200R_E_INVALID; }
    if (max_chunks > cap || n > rcap) { set_error("partitioned routing: batch (n=%lld, max_chunks=%lld) exceeds the exchange capacity (rcap=%lld, cap=%lld)", (long long)n, (long long)max_chunks, (long long)rcap, (long long)cap); return B200R_E_INVALID; }
    flags &= ~(uint32_t)(B200R_F_PREP_ONLY | B200R_F_PREP_DONE);
    const PartLayout L = part_layout(world, cap, rcap);
    const Plan p = make_plan(n, max_chunks, flags, true);
    uint8_t *sc = (uint8_t *)d_scratch, *wk = (uint8_t *)d_work;
    const int set = (int)(epoch % PART_SETS);
    unsigned long long *totals = (unsigned long long *)(wk + L.w_totals + (size_t)set * align_up((size_t)world * 8));
    uint32_t *dest_slot = (uint32_t *)(wk + L.w_dest_slot + (size_t)set * align_up((size_t)rcap * 4));
    const bool direct = (flags & B200R_F_ALL_ASCII) && n > 0 && !g_part_staged;
    if (n > 0) {
        rc = check_prompt_args(d_bytes, d_starts, d_lens, n, max_chunks, t->chunk_chars);
        if (rc) return rc;
        if (!d_scratch || scratch_bytes < p.total) { set_error("scratch too small: need %zu bytes, got %zu", p.total, scratch_bytes); return B200R_E_SCRATCH; }
        PartPlace pp = PartPlace();
        if (direct) {
            pp.world = world; pp.my_rank = rank; pp.totals = totals; pp.cap = cap; pp.rcap = rcap;
            pp.peer_bases = (const unsigned long long *)d_peer_bases;
            pp.hash_off = L.hash_off + (size_t)set * L.hash_set;
            pp.info_off = L.info_off + (size_t)set * L.info_set;
            pp.dest_slot = dest_slot;
            pp.out_ptr = (unsigned long long *)(wk + L.w_out_ptr + (size_t)set * align_up((size_t)rcap * 8));
        }
        int32_t *cb = (int32_t *)(sc + p.off_chunk_base);
        uint64_t *hs = (uint64_t *)(sc + p.off_hashes);
        rc = launch_hash(d_bytes, d_starts, d_lens, d_nchars, n, t->chunk_chars, flags, cb, hs, max_chunks,
                         sc, p, t->dev, prep_st, 1, direct ? &pp : nullptr);
        if (rc) return rc;
        if (prep_st != st) {
            B200R_CUDA(cudaEventRecord(prepped, prep_st));
            B200R_CUDA(cudaStreamWaitEvent(st, prepped, 0));
        }
        rc = launch_hash(d_bytes, d_starts, d_lens, d_nchars, n, t->chunk_chars, flags, cb, hs, max_chunks,
                         sc, p, t->dev, st, 2, direct ? &pp : nullptr);
        if (rc) return rc;
    }
    PartDyn *dyn = part_dyn(wk, L, set);
    ProfScope ps__(PK_PUSH, st);
    if (n > 0 && !direct) {
        const int32_t *chunk_base = (const int32_t *)(sc + p.off_chunk_base);
        const uint64_t *hashes = (const uint64_t *)(sc + p.off_hashes);
        const int nblk = (int)((n + PART_BLOCK - 1) / PART_BLOCK);
        g_launches += 2;
        uint64_t *hash_stage = (uint64_t *)(wk + L.w_hash_stage);
        unsigned long long *info_stage = (unsigned long long *)(wk + L.w_info_stage);
        k_part_place<<<nblk, PART_BLOCK, 0, st>>>(chunk_base, hashes, (int)n, world, totals, cap, rcap,
                                                  info_stage, dest_slot, hash_stage);
        B200R_LAUNCH_CHECK("k_part_place");
        k_part_bulk<<<148 * 2, 256, 0, st>>>(hash_stage, info_stage, totals, world, rank,
                                             (const unsigned long long *)d_peer_bases,
                                             L.hash_off + (size_t)set * L.hash_set,
                                             L.info_off + (size_t)set * L.info_set, cap, rcap,
                                             &dyn->ticket_bulk, set, (unsigned long long)epoch, (unsigned)n);
        B200R_LAUNCH_CHECK("k_part_bulk");
    } else {
        g_launches += 1;
        k_part_publish_in<<<1, P2P_MAX_WORLD, 0, st>>>((const unsigned long long *)d_peer_bases, world, rank,
                                                       set, epoch, (unsigned)n, n > 0 ? totals : nullptr);
        B200R_LAUNCH_CHECK("k_part_publish_in");
    }
    return B200R_OK;
}

int b200r_part_send(b200r_trie *t, const uint8_t *d_bytes, const int64_t *d_starts,
                    const int32_t *d_lens, const int32_t *d_nchars, int64_t n, int64_t max_chunks,
                    uint32_t flags, const uint64_t *d_peer_bases, int32_t world, int32_t rank,
                    int64_t cap, int64_t rcap, uint32_t epoch, void *d_work, void *d_scratch,
                    size_t scratch_bytes, void *stream)
{
    return part_send_impl(t, d_bytes, d_starts, d_lens, d_nchars, n, max_chunks, flags, d_peer_bases,
                          world, rank, cap, rcap, epoch, d_work, d_scratch, scratch_bytes,
                          (cudaStream_t)stream, nullptr, (cudaStream_t)stream);
}

// Owner side: wait for every source's send, build the CSR of the owned requests, route them
// (st_route); push each source's results back (st_ret, after `routed`).  One stream for both
// when st_ret == st_route.
static int part_route_impl(b200r_trie *t, uint64_t avail_mask, uint64_t seed,
                           const uint64_t *d_peer_bases, void *d_my_base, int32_t world, int32_t rank,
                           int64_t cap, int64_t rcap, uint32_t epoch, void *d_work,
                           cudaStream_t st, cudaStream_t st_ret, cudaEvent_t routed);

int b200r_part_route(b200r_trie *t, uint64_t avail_mask, uint64_t seed,
                     const uint64_t *d_peer_bases, void *d_my_base, int32_t world, int32_t rank,
                     int64_t cap, int64_t rcap, uint32_t epoch, void *d_work, void *stream)
{
    return part_route_impl(t, avail_mask, seed, d_peer_bases, d_my_base, world, rank, cap, rcap, epoch,
                           d_work, (cudaStream_t)stream, (cudaStream_t)stream, nullptr);
}

static int part_route_impl(b200r_trie *t, uint64_t avail_mask, uint64_t seed,
                           const uint64_t *d_peer_bases, void *d_my_base, int32_t world, int32_t rank,
                           int64_t cap, int64_t rcap, uint32_t epoch, void *d_work,
                           cudaStream_t st, cudaStream_t st_ret, cudaEvent_t routed)
{
    int rc = part_check(world, rank, cap, rcap, d_peer_bases, d_work);
    if (rc) return rc;
    if (!t || !d_my_base) { set_error("b200r_part_route: null pointer"); return B200R_E_INVALID; }
    const PartLayout L = part_layout(world, cap, rcap);
    const Plan p = make_plan(L.n_max, L.c_max, B200R_F_ALL_ASCII, true);
    const int set = (int)(epoch % PART_SETS);
    uint8_t *base = (uint8_t *)d_my_base, *wk = (uint8_t *)d_work, *sc = wk + L.w_plan;
    PartDyn *dyn = part_dyn(wk, L, set);
    uint32_t *pos_owned = (uint32_t *)(wk + L.w_pos_owned);
    int32_t *ep = (int32_t *)(wk + L.w_ep + (size_t)set * align_up((size_t)L.n_max * 4));
    int32_t *depth = (int32_t *)(wk + L.w_depth + (size_t)set * align_up((size_t)L.n_max * 4));
    uint64_t *matched = (uint64_t *)(wk + L.w_matched + (size_t)set * align_up((size_t)L.n_max * 8));
    int32_t *chunk_base = (int32_t *)(sc + p.off_chunk_base);
    uint32_t *slots = (uint32_t *)(sc + p.off_slots);
    int32_t *leader = (int32_t *)(sc + p.off_leader), *choice = (int32_t *)(sc + p.off_choice);
    const int32_t *n_dev = &dyn->n_owned;

    { ProfScope ps__(PK_WAIT, st); g_launches += 1;
    k_part_wait_meta<<<1, P2P_MAX_WORLD, 0, st>>>((const PartFlags *)base, set, (unsigned long long)epoch,
                                                  world, dyn, part_dyn(wk, L, (set + PART_SETS - 1) % PART_SETS), t->dev);
    B200R_LAUNCH_CHECK("k_part_wait_meta");
    }
    int32_t *chunk_end = (int32_t *)(wk + L.w_chunk_end);
    { ProfScope ps__(PK_PREP, st); g_launches += 1;
    k_part_gather<<<148, 256, 0, st>>>(base, L.info_off + (size_t)set * L.info_set, cap, rcap, world, dyn,
                                       chunk_base, chunk_end, pos_owned, (uint32_t *)(wk + L.w_idx_of_pos));
    B200R_LAUNCH_CHECK("k_part_gather");
    }
    // grids are estimates (about two shards' worth); the kernels stride over the rest
    RouteCore core = RouteCore();
    // the hashes are routed from where the sources stored them (inbound buffer of this set)
    core.hashes = (const uint64_t *)(base + L.hash_off + (size_t)set * L.hash_set);
    core.chunk_base = chunk_base; core.chunk_end = chunk_end;
    core.n = 2 * rcap < L.n_max ? 2 * rcap : L.n_max;
    core.max_chunks = L.c_max;
    core.avail = avail_mask; core.seed = seed; core.seq0 = 0;
    core.slots = slots; core.leader = leader; core.choice = choice;
    core.dpre = (int32_t *)(sc + p.off_dpre); core.cand = (uint64_t *)(sc + p.off_cand);
    core.endpoint = ep; core.depth = depth; core.matched = matched;
    core.dyn = &dyn->base; core.pos = pos_owned; core.n_dev = n_dev;
    core.idx_of_pos = (const uint32_t *)(wk + L.w_idx_of_pos);
    core.claim_kind = claim_kind_for(rcap, cap);
    rc = launch_route_core(t, core, st);
    if (rc) return rc;
    if (st_ret != st) {
        B200R_CUDA(cudaEventRecord(routed, st));
        B200R_CUDA(cudaStreamWaitEvent(st_ret, routed, 0));
    }
    { ProfScope ps__(PK_PUSH, st_ret); g_launches += 1;
    k_part_push_results<<<148 * 2, 256, 0, st_ret>>>(dyn, world, rank, ep, depth, matched,
                                                 (const unsigned long long *)d_peer_bases,
                                                 L.res_off + (size_t)set * L.res_set, rcap, set,
                                                 (unsigned long long)epoch);
    B200R_LAUNCH_CHECK("k_part_push_results");
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
    const int set = (int)(epoch % PART_SETS);
    uint8_t *base = (uint8_t *)d_my_base, *wk = (uint8_t *)d_work;
    { ProfScope ps__(PK_WAIT, st); g_launches += 1;
    k_part_wait_res<<<1, P2P_MAX_WORLD, 0, st>>>((const PartFlags *)base, set, (unsigned long long)epoch,
                                                 world, t->dev);
    B200R_LAUNCH_CHECK("k_part_wait_res");
    }
    if (n > 0) {
        ProfScope ps__(PK_RESOLVE, st); g_launches += 1;
        k_part_scatter<<<(int)((n + 255) / 256), 256, 0, st>>>(
            (const uint32_t *)(wk + L.w_dest_slot + (size_t)set * align_up((size_t)rcap * 4)), base,
            L.res_off + (size_t)set * L.res_set, rcap, (int)n, d_endpoint, d_match_depth, d_matched);
        B200R_LAUNCH_CHECK("k_part_scatter");
    }
    return B200R_OK;
}

// One call per batch: the three phases with the stream handling done natively (in Python it
// cost more host time per batch than the GPU needs).  Four internal streams:
//   prep  (high)  H2D copies of the inputs, per-request counting + scan of batch b
//   send  (low)   hashing kernel (fills every SM), placement, shipment to the owners
//   route (high)  wait for the sources, gather, claim, select, commit
//   ret   (high)  push results to the sources, wait for the owners' results, scatter
// so the route chain of batch b+1 starts right behind the commit of batch b, and the send of
// batch b+1 overlaps both.  `d_scratch` alternates between PART_SETS buffers, batch b: b % PART_SETS (the
// caller passes the one of this batch); dyn blocks, result buffers and exchange buffer sets
// rotate the same way.
static int part_streams(b200r_trie *t)
{
    if (t->part_ready) return B200R_OK;
    int least = 0, greatest = 0;
    B200R_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    B200R_CUDA(cudaStreamCreateWithPriority(&t->part_send, cudaStreamNonBlocking, least));
    B200R_CUDA(cudaStreamCreateWithPriority(&t->part_route, cudaStreamNonBlocking, greatest));
    B200R_CUDA(cudaStreamCreateWithPriority(&t->part_prep, cudaStreamNonBlocking, greatest));
    B200R_CUDA(cudaStreamCreateWithPriority(&t->part_ret, cudaStreamNonBlocking, greatest));
    cudaEvent_t *evs[] = { &t->part_inputs, &t->part_prepped, &t->part_routed };
    for (cudaEvent_t *e : evs) B200R_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    for (int k = 0; k < PART_SETS; k++) {
        B200R_CUDA(cudaEventCreateWithFlags(&t->part_sent[k], cudaEventDisableTiming));
        B200R_CUDA(cudaEventCreateWithFlags(&t->part_finished[k], cudaEventDisableTiming));
    }
    t->part_ready = true;
    return B200R_OK;
}

int b200r_part_begin(b200r_trie *t, uint32_t epoch)
{
    if (!t) { set_error("null trie"); return B200R_E_INVALID; }
    int rc = part_streams(t);
    if (rc) return rc;
    // Batch `epoch` reuses what batch epoch-3 used (PART_SETS buffer sets; the events of that set
    // still belong to batch epoch-3 here).  Its send phase overwrites that exchange buffer set: it may start
    // once this rank has finished batch epoch-3 (every owner has then pushed its results for
    // that batch, i.e. has consumed what all sources sent it).  Its prep writes the scratch
    // buffer of that set: after the send of batch epoch-3 has read it.
    const int par = (int)(epoch % PART_SETS);
    // One wait does it: the prep stream stores into that set first, the send stream follows the
    // prep stream (prepped event), and "returned" implies that the send of that batch had been
    // consumed by every owner, i.e. had read its scratch.
    if (t->part_has_finished[par])
        B200R_CUDA(cudaStreamWaitEvent(t->part_prep, t->part_finished[par], 0));
    return B200R_OK;
}

int b200r_part_copy_h2d(b200r_trie *t, void *d_dst, const void *h_src, size_t bytes)
{
    if (!t || !t->part_ready || (bytes && (!d_dst || !h_src))) { set_error("b200r_part_copy_h2d: bad argument (call b200r_part_begin first)"); return B200R_E_INVALID; }
    if (bytes) B200R_CUDA(cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, t->part_prep));
    return B200R_OK;
}

int b200r_part_step(b200r_trie *t, const uint8_t *d_bytes, const int64_t *d_starts,
                    const int32_t *d_lens, const int32_t *d_nchars, int64_t n, int64_t max_chunks,
                    uint32_t flags, uint64_t avail_mask, uint64_t seed, const uint64_t *d_peer_bases,
                    void *d_my_base, int32_t world, int32_t rank, int64_t cap, int64_t rcap,
                    uint32_t epoch, void *d_work, void *d_scratch, size_t scratch_bytes,
                    int32_t *d_endpoint, int32_t *d_match_depth, uint64_t *d_matched,
                    int32_t inputs_on_prep_stream, void *stream)
{
    int rc = b200r_part_begin(t, epoch);
    if (rc) return rc;
    cudaStream_t cur = (cudaStream_t)stream;
    const int me = (int)(epoch % PART_SETS);
    B200R_CUDA(cudaEventRecord(t->part_inputs, cur));
    if (!inputs_on_prep_stream) B200R_CUDA(cudaStreamWaitEvent(t->part_prep, t->part_inputs, 0));
    rc = part_send_impl(t, d_bytes, d_starts, d_lens, d_nchars, n, max_chunks, flags, d_peer_bases,
                        world, rank, cap, rcap, epoch, d_work, d_scratch, scratch_bytes,
                        t->part_prep, t->part_prepped, t->part_send);
    if (rc) return rc;
    B200R_CUDA(cudaEventRecord(t->part_sent[me], t->part_send));
    B200R_CUDA(cudaStreamWaitEvent(t->part_route, t->part_sent[me], 0));
    // (the route stage's dyn block and result buffers of this set are free: the send waited for
    // batch epoch-3 to be returned)
    rc = part_route_impl(t, avail_mask, seed, d_peer_bases, d_my_base, world, rank, cap, rcap, epoch,
                         d_work, t->part_route, t->part_ret, t->part_routed);
    if (rc) return rc;
    B200R_CUDA(cudaStreamWaitEvent(t->part_ret, t->part_inputs, 0));        // the outputs may be in use
    rc = b200r_part_finish(t, n, d_my_base, world, rank, cap, rcap, epoch, d_work, d_endpoint,
                           d_match_depth, d_matched, t->part_ret);
    if (rc) return rc;
    B200R_CUDA(cudaEventRecord(t->part_finished[me], t->part_ret));
    t->part_has_finished[me] = true;
    B200R_CUDA(cudaStreamWaitEvent(cur, t->part_finished[me], 0));
    return B200R_OK;
}

}  // extern "C"
