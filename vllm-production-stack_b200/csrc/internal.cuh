// internal.cuh -- host-side pieces shared by prefix.cu (hashing, single-GPU routing, C ABI)
// and multigpu.cu (the partitioned multi-GPU exchange).
#pragma once
#include <vector>

#include "trie.cuh"

namespace b200r {

// ------------------------------------------------------------- profiling
// Optional per-kernel timing with CUDA events on the launching stream
// (b200r_prof_enable / b200r_prof_read).  Off by default: the hot path then
// records nothing.  g_launches counts every kernel this library enqueues.
enum { PK_PREP = 0, PK_UTF8, PK_HASH, PK_CLAIM, PK_SELECT, PK_COMMIT, PK_LOOKUP, PK_INSERT,
       PK_CLASSIFY, PK_PUSH, PK_WAIT, PK_APPLY, PK_RESOLVE, PK_COUNT };
struct ProfSpan { int kid; cudaEvent_t a, b; };
extern bool g_prof_on;
extern std::vector<ProfSpan> g_spans;
extern long long g_launches;
extern int g_claim_mode;
extern int g_hash_mode;
extern int g_part_staged;      // b200r_set_part_staged: 1 = always compact + ship (test / comparison)
int claim_kind_for(int64_t n, int64_t chunks);

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

constexpr int WARPS_PER_CTA = 8;       // per-request kernels: one warp per request
constexpr int WALK_MAX_CTAS = 148 * 8; // persistent grid of the walk kernels (<= NWSLOT warps)

// Device-resident bookkeeping of the partitioned multi-GPU path (read by k_select).
constexpr int P2P_MAX_WORLD = 64;
constexpr int PART_SETS = 3;       // buffer sets of the partitioned exchange: batch b uses set b % 3
                                   // and may start once batch b - 3 has been returned
struct P2PDyn {
    unsigned long long seq_base;     // global sequence number of the next batch's request 0
    uint32_t my_off, total_n;        // position of this shard / size of the current global batch
};

static inline size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }

// Partitioned multi-GPU path, send side: the prep kernel also decides where every request goes
// (owner = mix(first chunk hash) % world, slot in the (source, owner) segment taken with one
// packed atomicAdd {requests << 32 | chunks} per warp and owner), writes the request's info
// word straight into the owner's inbound buffer and leaves, per request, the ADDRESS of its
// first chunk hash in that buffer -- where the hashing kernel then stores the hashes (plain
// stores over NVLink peer memory; no staging copy, no shipment kernel).  world == 0: off.
struct PartPlace {
    int world, my_rank;
    unsigned long long *totals;             // [world]
    long long cap, rcap;
    const unsigned long long *peer_bases;   // [world]
    size_t hash_off, info_off;              // of this batch's buffer set
    uint32_t *dest_slot;                    // [n] owner << 24 | slot
    unsigned long long *out_ptr;            // [n]
};

__host__ __device__ inline int part_owner_of(uint64_t h0, int world)
{
    const uint64_t x = h0 * 0x9E3779B97F4A7C15ULL;
    return (int)((x >> 33) % (uint64_t)world);
}

struct Plan {           // carve-up of the caller's scratch buffer
    size_t off_tile_first, off_block_sums, off_nchars, off_chunk_off, off_chunk_base, off_hashes,
        off_slots, off_leader, off_choice, off_dpre, off_cand, off_res, total;
};
Plan make_plan(int64_t n, int64_t max_chunks, uint32_t flags, bool routing);

int check_prompt_args(const void *bytes, const void *starts, const void *lens, int64_t n,
                      int64_t max_chunks, int32_t cc);

// prep + (utf8) + chunk hashing
int launch_hash(const uint8_t *d_bytes, const int64_t *d_starts, const int32_t *d_lens,
                const int32_t *d_nchars, int64_t n, int32_t cc, uint32_t flags,
                int32_t *d_chunk_base, uint64_t *d_hashes, int64_t max_chunks, uint8_t *scratch,
                const Plan &p, TrieDev *dev, cudaStream_t st, int phase = 0,
                const PartPlace *pp = nullptr);

// claim -> select -> commit over chunk hashes already in memory.  `n` sizes the grids; when the
// real count lives on the device (n_dev) it is an estimate and the kernels stride.
struct RouteCore {
    const uint64_t *hashes; const int32_t *chunk_base; const int32_t *chunk_end;  // NULL = CSR
    int64_t n; int64_t max_chunks;
    uint64_t avail, seed, seq0;
    uint32_t *slots; int32_t *leader, *choice, *dpre; uint64_t *cand;
    int32_t *endpoint, *depth; uint64_t *matched;
    const P2PDyn *dyn; const uint32_t *pos; const uint32_t *idx_of_pos; const int32_t *n_dev;
    unsigned long long *res;   // per request, tagged {root, endpoint | leader}: the fused select + commit
                               // kernel resolves leader chains through it (NULL: two kernels)
    int claim_kind;       // claim_kind_for()
};
int launch_route_core(b200r_trie *t, const RouteCore &rc, cudaStream_t st);

TrieRef trie_ref(const b200r_trie *t);

}  // namespace b200r

struct b200r_trie {
    int device;
    int cap_log2;
    int chunk_chars;
    b200r::Head *tab;
    b200r::Rec *log;
    uint2 *wblk;
    b200r::TrieDev *dev;
    uint32_t epoch;
    // streams / events of b200r_part_step (created on first use)
    cudaStream_t part_send, part_route, part_prep, part_ret;
    cudaEvent_t part_inputs, part_prepped, part_routed, part_sent[b200r::PART_SETS], part_finished[b200r::PART_SETS];
    bool part_ready, part_has_finished[b200r::PART_SETS];
    // latency path (b200r_prefix_route_one; allocated on first use)
    uint8_t *one_host, *one_host_dev, *one_dev;
};
