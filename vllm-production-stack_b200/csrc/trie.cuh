// trie.cuh -- the GPU-resident HashTrie: a node log plus a hash table of run heads.
//
// Reference structure: TrieNode {children: dict[chunk hash -> TrieNode], endpoints: set}
// (src/vllm_router/prefix/hashtrie.py:24-32).  Here every trie node is one 32-byte record
// in an append-only LOG, and a node's identity is its log index.  The nodes a request
// creates in one go (the part of its path that did not exist yet) are laid out as one
// contiguous RUN of records, in depth order, so that
//   - creating a path is one coalesced write of its records,
//   - walking along a stored path is a coalesced read of consecutive records whose chunk
//     hashes are compared with the request's (no hashing, no random access), and
//   - only the FIRST node of a run has to be findable: the open-addressing hash table maps
//     (parent node index, chunk hash) -> (log index, run length) for run heads only.  That
//     key is exactly `parent.children[chunk_hash]`, so node identity is exact -- no path
//     fingerprints and no collision probability beyond XXH64's own.
// A node is therefore either a run head (in the table) or the in-run successor of its
// parent (the record right behind the parent's record).  Every node still has its own
// endpoint mask and batch bookkeeping word, so the routing semantics are node-per-chunk
// exactly like the reference; only the addressing differs.  A request that shares a stored
// path up to some depth and then diverges starts a new run whose head hangs under a node
// in the middle of the older run.
#pragma once
#include "common.cuh"

namespace b200r {

struct __align__(32) Rec {       // one trie node
    uint64_t h;                  // chunk hash of the edge parent -> node
    uint64_t rc;                 // low 32 bits: records from this one to the end of its run (>= 1);
                                 // high 32 bits: check word of h (rec_chk)
    uint64_t mask;               // endpoint set                     } only ever changed
    unsigned long long first;    // (epoch << 32) | ~request: first   } by atomics
                                 // request of batch `epoch` that touched the node (atomicMax); 0 = none
};
// Publication needs no fence.  The log starts zeroed; a record is written exactly once: {h, rc}
// by one 16-byte store, while mask and first start at zero and are only ever updated by
// atomics (the creator's own first-toucher mark included), so nothing a reader does can be
// overwritten by a store that lands late.  A run becomes reachable when its head entry appears
// in the table (one 128-bit CAS carrying key and log index); its records may become visible a
// little later: a reader that finds a record of a published run with rc == 0, or whose check
// word does not fit its h (the two words of the store seen at different moments), re-reads it
// past L1 until it is whole.

struct __align__(16) Head {      // hash-table entry: the run that starts at child `h` of `parent`
    uint64_t h;                  // chunk hash of the run's first node
    uint64_t pk;                 // (parent key << 32) | log index of the first node; 0 = empty
};                               // parent key = parent's log index + 2, the root's is 1

struct TrieDev {                 // device-resident control block
    uint64_t root_mask;
    unsigned long long num_nodes;    // filled on demand (status / dump), not on the hot path
    unsigned int status;
    unsigned int pad0;
    unsigned long long dump_count;
    unsigned long long log_count;    // bump allocator of the node log (may overshoot log_cap)
};

struct TrieRef {                 // by-value handle for kernels
    Head *tab;                   // 2^tab_log2 entries = 2 per node of capacity: load <= 0.5
    Rec *log;                    // log_cap = 2^cap_log2 records
    TrieDev *dev;
    uint2 *wblk;                 // per warp slot: private allocation block [x, y) kept across launches
    int tab_log2;
    uint32_t log_cap;
};

constexpr uint32_t ROOT_IDX = 0xFFFFFFFFu;      // "log index" of the root (it has no record)
constexpr uint32_t SLOT_NONE = 0xFFFFFFFFu;
constexpr uint32_t PREOK_BIT = 0x80000000u;     // slots[]: node usable before the batch
constexpr uint32_t IDX_MASK = 0x7FFFFFFFu;
constexpr int NWSLOT = 65536;                   // tile slots with a persistent allocation block
constexpr unsigned FULLW = 0xffffffffu;
#define B200R_S_INTERNAL 8u                     /* latched: inconsistent table (never expected) */

__device__ __forceinline__ uint32_t parent_key(uint32_t pidx) { return pidx + 2u; }   // ROOT_IDX -> 1

// ------------------------------------------------------------------ memory ops
__device__ __forceinline__ uint32_t rec_chk(uint64_t h)
{
    return ((uint32_t)h * 0x9E3779B1u) ^ ((uint32_t)(h >> 32) * 0x85EBCA77u) ^ 0x5bd1e995u;
}
__device__ __forceinline__ bool rec_whole(uint64_t h, uint64_t rc)
{
    return (uint32_t)rc != 0 && (uint32_t)(rc >> 32) == rec_chk(h);
}
__device__ __forceinline__ void ld_rec(const Rec *r, uint64_t &h, uint64_t &mask, unsigned long long &first,
                                       uint64_t &rc)
{   // one 32-byte load (LDG.256).  L1 may serve it: h / rc never change once written, masks only
    // change in kernels that do not read them through this path, and a stale `first` only
    // costs a redundant atomicMax.
    asm volatile("ld.global.ca.v4.u64 {%0, %1, %2, %3}, [%4];"
                 : "=l"(h), "=l"(rc), "=l"(mask), "=l"(first) : "l"(r) : "memory");
}
__device__ __forceinline__ void ld_rec_cg(const Rec *r, uint64_t &h, uint64_t &mask, unsigned long long &first,
                                          uint64_t &rc)
{
    asm volatile("ld.global.cg.v4.u64 {%0, %1, %2, %3}, [%4];"
                 : "=l"(h), "=l"(rc), "=l"(mask), "=l"(first) : "l"(r) : "memory");
}
__device__ __forceinline__ void st_rec_key(Rec *r, uint64_t h, uint32_t rem)
{   // the immutable half of a record
    const uint64_t rc = (uint64_t)rem | ((uint64_t)rec_chk(h) << 32);
    asm volatile("st.global.v2.u64 [%0], {%1, %2};" :: "l"(r), "l"(h), "l"(rc) : "memory");
}
__device__ __forceinline__ void ld_head(const Head *e, uint64_t &a, uint64_t &b)
{
    asm volatile("ld.global.ca.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(e) : "memory");
}
__device__ __forceinline__ void ld_head_cg(const Head *e, uint64_t &a, uint64_t &b)
{
    asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(e) : "memory");
}
template <bool RELEASE = false>
__device__ __forceinline__ void cas128(Head *e, uint64_t n1, uint64_t n2, uint64_t &o1, uint64_t &o2)
{   // compare the entry with (0,0), swap in (n1,n2); returns the previous entry.  RELEASE: every
    // write that happens-before this thread's CAS is visible to whoever sees the new entry.
    if (RELEASE) {
        asm volatile(
            "{\n\t"
            ".reg .b128 cmp, nw, old;\n\t"
            "mov.b128 cmp, {%2, %3};\n\t"
            "mov.b128 nw, {%4, %5};\n\t"
            "atom.global.release.gpu.cas.b128 old, [%6], cmp, nw;\n\t"
            "mov.b128 {%0, %1}, old;\n\t"
            "}\n"
            : "=l"(o1), "=l"(o2)
            : "l"(0ULL), "l"(0ULL), "l"(n1), "l"(n2), "l"(e)
            : "memory");
        return;
    }
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

__device__ __forceinline__ uint32_t slot_of(uint64_t h, uint32_t pkey, int tab_log2)
{   // 32-bit mixing (h is a hash already): three multiplies on the 32-bit pipe
    uint32_t x = ((uint32_t)h ^ __funnelshift_l((uint32_t)(h >> 32), (uint32_t)(h >> 32), 15)) * 0x9E3779B1u;
    x += pkey * 0x85EBCA77u;
    x ^= x >> 15;
    x *= 0x2C1B3C6Du;
    return x >> (32 - tab_log2);
}

// ------------------------------------------------------------------ head table
// Log index of child `h` of the node with parent key `pkey`, if it heads a run; else SLOT_NONE.
// An "absent" answer may come from a stale L1 line; callers that act on it (creation) confirm
// with the CAS in head_publish.  coherent = read past L1 from the start.
__device__ __forceinline__ uint32_t head_find(const Head *tab, int tab_log2, uint64_t h, uint32_t pkey,
                                              bool coherent, uint32_t &eslot)
{   // eslot: the empty slot the probe sequence ended at when the answer is "absent"
    const uint32_t m = (1u << tab_log2) - 1;
    uint32_t s = slot_of(h, pkey, tab_log2);
    eslot = SLOT_NONE;
    for (uint32_t probe = 0; probe <= m; probe++) {
        uint64_t a, b;
        if (coherent) ld_head_cg(tab + s, a, b);
        else ld_head(tab + s, a, b);
        if (b == 0) { eslot = s; return SLOT_NONE; }            // empty slot: not there
        if (a == h && (uint32_t)(b >> 32) == pkey) return (uint32_t)b;
        s = (s + 1) & m;
    }
    return SLOT_NONE;
}

// Enter the run starting at log index idx under (h, pkey).  Returns idx when this call entered
// it; the log index of the existing run when that child exists already (someone else created
// it first); SLOT_NONE if the table is full.
template <bool RELEASE = false>
__device__ __forceinline__ uint32_t head_publish(Head *tab, int tab_log2, uint64_t h, uint32_t pkey,
                                                 uint32_t idx)
{
    const uint32_t m = (1u << tab_log2) - 1;
    const uint64_t pk = ((uint64_t)pkey << 32) | idx;
    uint32_t s = slot_of(h, pkey, tab_log2);
    for (uint32_t probe = 0; probe <= m; probe++) {
        Head *e = tab + s;
        uint64_t a, b;
        ld_head(e, a, b);
        if (b == 0) {
            cas128<RELEASE>(e, h, pk, a, b);
            if (a == 0 && b == 0) return idx;
        }                                       // else (a, b) is what sits there now
        if (a == h && (uint32_t)(b >> 32) == pkey) return (uint32_t)b;
        s = (s + 1) & m;
    }
    return SLOT_NONE;
}

// ------------------------------------------------------------------ log allocation
// A warp keeps a private block of log records [cur, end) in registers and, between launches,
// in t.wblk[warp slot], so a run costs no global atomic in the common case.  Runs at least as
// long as the block size are allocated directly.  The block size scales with the log so that
// the space parked in blocks stays a small part of a small table.
struct WarpAlloc { uint32_t cur, end; };

__device__ __forceinline__ uint32_t alloc_block_size(uint32_t log_cap)
{
    const uint32_t b = log_cap >> 17;
    return b > 128u ? 128u : b;
}

// A request is walked by a TILE of TW lanes (8, 16 or the whole warp): several short requests
// share a warp, each tile on its own.  Collectives use the tile's lane mask.
struct Tile { unsigned mask; int base; int tl; };       // lanes of the tile, first lane, my index
template <int TW>
__device__ __forceinline__ Tile make_tile(int lane)
{
    Tile g;
    g.base = lane & ~(TW - 1);
    g.tl = lane & (TW - 1);
    g.mask = TW == 32 ? FULLW : (((1u << TW) - 1u) << g.base);
    return g;
}
template <int TW>
__device__ __forceinline__ unsigned tile_ballot(const Tile &g, bool p)
{   // bit k = lane k of the tile
    const unsigned b = __ballot_sync(g.mask, p);
    return TW == 32 ? b : ((b >> g.base) & ((1u << TW) - 1u));
}
template <int TW>
__device__ __forceinline__ int leading_ones_t(unsigned bal)
{   // bal has no bits at or above TW
    return (TW == 32 && bal == FULLW) ? 32 : (__ffs(~bal) - 1);
}

// tile-uniform: every lane passes the same arguments and receives the same result
__device__ __forceinline__ uint32_t warp_alloc(const TrieRef &t, WarpAlloc &wa, uint32_t L, const Tile &g)
{
    const int lane = g.tl;
    if (L <= wa.end - wa.cur) {
        const uint32_t base = wa.cur;
        wa.cur += L;
        return base;
    }
    const uint32_t blk = alloc_block_size(t.log_cap);
    const uint32_t take = L >= blk ? L : blk;
    unsigned long long b = 0;
    if (lane == 0) b = atomicAdd(&t.dev->log_count, (unsigned long long)take);
    b = __shfl_sync(g.mask, b, g.base);
    if (b + L > (unsigned long long)t.log_cap) {
        if (lane == 0) atomicOr(&t.dev->status, B200R_S_TABLE_FULL);
        return SLOT_NONE;
    }
    if (take > L) {                             // a fresh block; what is left of the old one is dropped
        unsigned long long e = b + take;
        if (e > (unsigned long long)t.log_cap) e = t.log_cap;
        wa.cur = (uint32_t)b + L;
        wa.end = (uint32_t)e;
    }
    return (uint32_t)b;
}

// ------------------------------------------------------------------ request walk
struct ReqSpan { int b0, C; };

// chunk_end == NULL: plain CSR (request i ends where i + 1 begins); else the requests' hash
// runs are not adjacent (partitioned multi-GPU path: they lie where the sources stored them)
__device__ __forceinline__ ReqSpan req_span(const int32_t *chunk_base, const int32_t *chunk_end, int i,
                                            int64_t max_chunks)
{
    ReqSpan s;
    s.b0 = chunk_base[i];
    int64_t e = chunk_end ? chunk_end[i] : chunk_base[i + 1];
    if (e > max_chunks) e = max_chunks;
    s.C = (e > s.b0) ? (int)(e - s.b0) : 0;
    return s;
}

enum { W_LOOKUP = 0, W_CLAIM = 1, W_INSERT = 2 };

struct WalkArgs {
    TrieRef t;
    const uint64_t *hashes;
    const int32_t *chunk_base;
    const int32_t *chunk_end;     // NULL = CSR (see req_span)
    int64_t max_chunks;
    uint64_t avail;               // lookup / claim
    unsigned long long tag;       // claim: epoch << 32
    uint32_t *slots;              // claim: log index (| PREOK_BIT) per chunk
    int32_t *dpre;                // claim: per request, number of leading nodes usable before the batch
    uint64_t *cand;               // claim: per request, mask & avail of the deepest such node (avail if none)
    int32_t *depth_out;           // lookup
    uint64_t *matched_out;        // lookup
    const int32_t *endpoint;      // insert
    unsigned long long *res;      // claim (thread kernel): per-request word of the fused select +
                                  // commit kernel, zeroed here (scratch may hold another table's words)
    const uint32_t *pos;          // claim: global position of request i (NULL: i itself) -- the
                                  // index that "first toucher" marks are expressed in
};

__device__ __forceinline__ int leading_ones(unsigned bal)
{
    return (bal == FULLW) ? 32 : (__ffs(~bal) - 1);
}

// One warp walks request i down the trie, 32 depths (one window) per step.
//   B  the child of the last resolved node, if it exists and is not the record right behind
//      it, heads a run: one lane probes the hash table with (parent index, chunk hash).
//   A  from a known log position the next nodes can only be consecutive records: lanes read
//      records base+k (coalesced 32-byte loads; the first one tells how long the run is)
//      and compare their chunk hashes with their own chunks; the leading matches are resolved.
//      A mismatch or the end of the run sends the walk back to B.
//   absent  no record and no head: by prefix closure nothing deeper exists either.
//      W_CLAIM / W_INSERT append the rest of the path as ONE run and enter its head in the
//      table with a 128-bit CAS.  If another request of the batch entered that child first, the
//      walk simply goes on through the winner's run (the records just written stay unused).
struct Pending {              // a head CAS whose outcome has not been looked at yet
    int req;                   // request it belongs to, -1 = none
    int lane;                  // lane that issued it
    uint64_t oa, ob;           // what the CAS found in the slot ((0,0) = the run is entered)
};

// DEFER (claim fast path): the CAS that enters a new run goes straight to the empty slot the
// probe ended at and its result is left in `pend`; the caller looks at it one request later
// (the round trip overlaps the next request) and walks the request again, without DEFER, if
// the slot had been taken meanwhile.  Walking a request twice is harmless: every step is
// idempotent, and the records written the first time simply stay unused.
template <int MODE, bool DEFER, int TW>
__device__ __forceinline__ void walk_request(const WalkArgs &a, int i, const Tile &g, WarpAlloc &wa,
                                             const ReqSpan sp, uint64_t h_first, Pending &pend)
{
    const int lane = g.tl;                // position inside the tile
    const int C = sp.C;
    const uint64_t *hs = a.hashes + sp.b0;
    const TrieRef &t = a.t;
    uint64_t bit = 0;
    if (MODE == W_INSERT) {
        const int e = a.endpoint[i];
        if (e < 0 || e > 63) return;
        bit = 1ULL << e;
        if (lane == 0 && !(t.dev->root_mask & bit)) atomicOr((unsigned long long *)&t.dev->root_mask, bit);
    }
    const unsigned long long mine = a.tag | (unsigned long long)(~(a.pos ? a.pos[i] : (unsigned)i));
    uint32_t pidx = ROOT_IDX;             // last resolved node
    uint32_t base = 0, room = 0;          // A: candidates start at log[base]; room = records of the
    bool have_base = false;               //    run from base on (0 = not known yet)
    bool lost = !DEFER && MODE == W_CLAIM;  // second walk of a request: probe the table past L1
    int npre = 0;                         // leading nodes with mask & avail != 0 (lookup depth / claim dpre)
    bool allpre = true;
    uint64_t sel = a.avail;
    bool finished = false;

    for (int w0 = 0; w0 < C && !finished; w0 += TW) {
        const int d = w0 + lane;
        const bool valid = d < C;
        const int nvalid = (C - w0) < TW ? (C - w0) : TW;
        const uint64_t h = w0 == 0 ? h_first : (valid ? hs[d] : 0);
        int r = 0;                        // lanes [0, r) of the window are resolved
        while (r < nvalid) {
            const int off = lane - r;
            if (!have_base) {
                // ---- B
                uint32_t idx = SLOT_NONE, eslot = SLOT_NONE;
                if (off == 0) idx = head_find(t.tab, t.tab_log2, h, parent_key(pidx), lost, eslot);
                idx = __shfl_sync(g.mask, idx, g.base + r);
                if (idx == SLOT_NONE) {
                    // ---- absent at depth w0 + r, and so is everything below it
                    if (MODE == W_LOOKUP) { finished = true; break; }
                    const int pos = w0 + r;
                    const uint32_t L = (uint32_t)(C - pos);
                    const uint32_t nb = warp_alloc(t, wa, L, g);
                    if (nb == SLOT_NONE) {
                        if (MODE == W_CLAIM)
                            for (int d2 = d; d2 < C; d2 += TW)
                                if (d2 >= pos) a.slots[sp.b0 + d2] = SLOT_NONE;
                        finished = true;
                        break;
                    }
                    // every lane keeps to its own depths d, d + TW, ...: a slot word is only ever
                    // written by one thread, so a later correction (lost race) lands in order
                    for (int d2 = d; d2 < C; d2 += TW) {
                        if (d2 < pos) continue;
                        const uint32_t x = nb + (uint32_t)(d2 - pos);
                        st_rec_key(t.log + x, d2 == d ? h : hs[d2], (uint32_t)(C - d2));
                        if (MODE == W_CLAIM) {
                            atomicMax(&t.log[x].first, mine);
                            a.slots[sp.b0 + d2] = x;
                        } else {
                            atomicOr((unsigned long long *)&t.log[x].mask, bit);
                        }
                    }
                    if (DEFER) {                    // enter the run at the empty slot; look later
                        pend.req = i;
                        pend.lane = r;
                        pend.oa = 0; pend.ob = 1;   // "taken" unless the CAS below says otherwise
                        if (off == 0 && eslot != SLOT_NONE)
                            cas128(t.tab + eslot, h, ((uint64_t)parent_key(pidx) << 32) | nb, pend.oa, pend.ob);
                        finished = true;
                        break;
                    }
                    uint32_t got = 0;
                    if (off == 0) got = head_publish(t.tab, t.tab_log2, h, parent_key(pidx), nb);
                    got = __shfl_sync(g.mask, got, g.base + r);
                    if (got == nb) { finished = true; break; }          // the path is complete
                    if (got == SLOT_NONE) {
                        if (lane == 0) atomicOr(&t.dev->status, B200R_S_TABLE_FULL);
                        finished = true;
                        break;
                    }
                    lost = true;                                        // walk on through the winner's run
                    idx = got;
                }
                base = idx;
                room = 0;
                have_base = true;
            }
            // ---- A
            uint32_t idx = base + (uint32_t)off;
            uint64_t rh = 0, m = 0, rc = 0;
            unsigned long long f = 0;
            const uint32_t room0 = room;
            const bool rd = valid && off >= 0 && (room ? (uint32_t)off < room : idx < t.log_cap);
            if (rd) ld_rec(t.log + idx, rh, m, f, rc);
            if (room == 0) {                                    // run length from its first record
                if (off == 0) while (!rec_whole(rh, rc)) ld_rec_cg(t.log + idx, rh, m, f, rc);
                room = (uint32_t)__shfl_sync(g.mask, rc, g.base + r);
            }
            const bool in = valid && off >= 0 && (uint32_t)off < room;
            if (in) while (!rec_whole(rh, rc)) ld_rec_cg(t.log + idx, rh, m, f, rc);   // published, not yet visible
            const bool fresh = room0 == 0;                      // base came from the table just now
            const int nres = leading_ones_t<TW>(tile_ballot<TW>(g, in && rh == h) >> r);
            if (nres == 0) {
                if (fresh) {                                    // a head always matches itself
                    if (lane == 0) atomicOr(&t.dev->status, B200R_S_INTERNAL);
                    finished = true;
                    break;
                }
                have_base = false;                              // the path leaves the run right here
                continue;
            }
            pidx = base + (uint32_t)nres - 1;
            if (r + nres < TW || (uint32_t)nres == room) have_base = false;     // diverged, or the run ended
            else { base += (uint32_t)nres; room -= (uint32_t)nres; }            // next window, same run
            // ---- lanes [r, r + nres) hold resolved nodes
            const bool res = off >= 0 && off < nres;
            const bool use = res && (m & a.avail);
            if (MODE == W_CLAIM) {
                if (res) {
                    if (use) {
                        a.slots[sp.b0 + d] = idx | PREOK_BIT;
                    } else {
                        a.slots[sp.b0 + d] = idx;
                        if (f < mine) atomicMax(&t.log[idx].first, mine);
                    }
                }
            } else if (MODE == W_INSERT) {
                if (res && !(m & bit)) atomicOr((unsigned long long *)&t.log[idx].mask, bit);
            }
            if (MODE != W_INSERT && allpre) {
                const int nok0 = leading_ones_t<TW>(tile_ballot<TW>(g, use) >> r);
                const int nok = nok0 < nres ? nok0 : nres;
                if (nok > 0) {
                    sel = __shfl_sync(g.mask, m & a.avail, g.base + r + nok - 1);
                    npre += nok;
                }
                if (nok < nres) {
                    allpre = false;
                    if (MODE == W_LOOKUP) { finished = true; break; }
                }
            }
            r += nres;
        }
    }
    if (lane == 0) {
        if (MODE == W_LOOKUP) {
            a.depth_out[i] = npre;
            a.matched_out[i] = sel;
        } else if (MODE == W_CLAIM) {
            a.dpre[i] = npre;
            a.cand[i] = sel;
        }
    }
}

// One tile of TW lanes per request, persistent grid-stride; the allocation block of each tile
// slot survives the launch in t.wblk.  The span of the request after next and the chunk hashes
// of the next one are fetched while the current request is walked.
template <int MODE, int TW>
__global__ void __launch_bounds__(256, 4)
k_walk(WalkArgs a, int n, const int32_t *__restrict__ n_dev)
{
    const Tile g = make_tile<TW>(threadIdx.x & 31);
    const int gt = (blockIdx.x * blockDim.x + threadIdx.x) / TW;       // global tile
    const int nt = (gridDim.x * blockDim.x) / TW;
    if (n_dev) n = *n_dev;
    WarpAlloc wa = { 0, 0 };
    constexpr bool ALLOC = (MODE == W_CLAIM || MODE == W_INSERT);
    constexpr bool DEFER = (MODE == W_CLAIM);
    if (ALLOC) { const uint2 v = a.t.wblk[gt]; wa.cur = v.x; wa.end = v.y; }
    Pending pend = { -1, 0, 0, 0 }, none = { -1, 0, 0, 0 };
    ReqSpan sp0 = { 0, 0 }, sp1 = { 0, 0 };
    uint64_t h0 = 0;
    if (gt < n) {
        sp0 = req_span(a.chunk_base, a.chunk_end, gt, a.max_chunks);
        if (g.tl < sp0.C) h0 = a.hashes[sp0.b0 + g.tl];
    }
    if (gt + nt < n) sp1 = req_span(a.chunk_base, a.chunk_end, gt + nt, a.max_chunks);
    for (int i = gt; i < n; i += nt) {
        ReqSpan sp2 = { 0, 0 };
        if (i + 2 * nt < n) sp2 = req_span(a.chunk_base, a.chunk_end, i + 2 * nt, a.max_chunks);
        uint64_t h1 = 0;
        if (i + nt < n && g.tl < sp1.C) h1 = a.hashes[sp1.b0 + g.tl];
        Pending mine = { -1, 0, 0, 0 };
        walk_request<MODE, DEFER, TW>(a, i, g, wa, sp0, h0, mine);
        if (DEFER) {
            if (pend.req >= 0 && __ballot_sync(g.mask, g.tl == pend.lane && (pend.oa | pend.ob))) {
                // the slot was taken while the CAS was on its way: walk that request again
                const ReqSpan spr = req_span(a.chunk_base, a.chunk_end, pend.req, a.max_chunks);
                walk_request<MODE, false, TW>(a, pend.req, g, wa, spr,
                                              g.tl < spr.C ? a.hashes[spr.b0 + g.tl] : 0, none);
            }
            pend = mine;
        }
        sp0 = sp1; sp1 = sp2; h0 = h1;
    }
    if (DEFER && pend.req >= 0 && __ballot_sync(g.mask, g.tl == pend.lane && (pend.oa | pend.ob))) {
        const ReqSpan spr = req_span(a.chunk_base, a.chunk_end, pend.req, a.max_chunks);
        walk_request<MODE, false, TW>(a, pend.req, g, wa, spr,
                                      g.tl < spr.C ? a.hashes[spr.b0 + g.tl] : 0, none);
    }
    if (ALLOC && g.tl == 0) a.t.wblk[gt] = make_uint2(wa.cur, wa.end);
}

// ------------------------------------------------------------------ claim, one THREAD per request
// The same walk as walk_request<W_CLAIM>, but a thread per request instead of a warp.  With
// runs, walking a stored prefix is a short loop over consecutive records that are usually in L1
// (the shared prefixes of a batch are the hot ones), one table probe per run, and at most one
// probe that misses; a thread does that in a few hundred instructions, so 32 requests cost a
// warp about as many issue slots as ONE request costs the warp-per-request kernel.  Accesses
// are per-thread 32-byte sectors rather than coalesced lines, which is what the records are.
// Right for batches of short prompts (tens of chunks); the warp kernel is kept for long ones.
// Log space for the new runs of a warp's 32 requests is taken with one atomic per warp.
static __global__ void __launch_bounds__(128)
k_claim_thread(WalkArgs a, int n, const int32_t *__restrict__ n_dev)
{
    const TrieRef &t = a.t;
    const int lane = threadIdx.x & 31;
    const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nw = (gridDim.x * blockDim.x) >> 5;
    if (n_dev) n = *n_dev;
    for (int wbase = gw * 32; wbase < n; wbase += nw * 32) {
        const int i = wbase + lane;
        const bool active = i < n;
        ReqSpan sp = { 0, 0 };
        if (active) sp = req_span(a.chunk_base, a.chunk_end, i, a.max_chunks);
        if (active && a.res) a.res[i] = 0;
        const int C = sp.C;
        const uint64_t *hs = a.hashes + sp.b0;
        uint32_t *sl = a.slots + sp.b0;
        const unsigned long long mine = a.tag | (unsigned long long)(~(a.pos ? a.pos[i] : (unsigned)i));
        int d = 0;                            // depths [0, d) are resolved
        uint32_t pidx = ROOT_IDX;             // last resolved node
        int npre = 0;
        bool allpre = true;
        uint64_t sel = a.avail;
        bool todo = active && C > 0;
        bool coherent = false;
        for (;;) {
            uint32_t eslot = SLOT_NONE;
            uint64_t hd = 0;                  // chunk hash at depth d when the walk stops there
            // ---- find: follow stored runs as far as they go
            while (todo && d < C) {
                uint64_t h = hs[d];
                uint32_t idx = head_find(t.tab, t.tab_log2, h, parent_key(pidx), coherent, eslot);
                // a long stretch is about to be created: look once more, past L1 -- when many
                // requests meet a new shared prefix at the same moment, most of them then find the
                // winner's run instead of writing (and wasting) a copy of their own
                if (idx == SLOT_NONE && !coherent && C - d >= 8)
                    idx = head_find(t.tab, t.tab_log2, h, parent_key(pidx), true, eslot);
                if (idx == SLOT_NONE) { hd = h; break; }
                // the head record first: it matches by its key and tells how long the run is
                uint32_t room, k = 1;
                {
                    uint64_t rh, m, rc;
                    unsigned long long f;
                    ld_rec(t.log + idx, rh, m, f, rc);
                    while (!rec_whole(rh, rc)) ld_rec_cg(t.log + idx, rh, m, f, rc);
                    room = (uint32_t)rc;
                    if (m & a.avail) {
                        if (allpre) { npre++; sel = m & a.avail; }
                    } else {
                        allpre = false;
                        sl[d] = idx;
                        if (f < mine) atomicMax(&t.log[idx].first, mine);
                    }
                    d++;
                }
                // then the records behind it, four at a time (independent loads in flight)
                bool left = false;
                while (!left && d < C && k < room) {
                    uint64_t rh[4], m[4], rc[4], hh[4];
                    unsigned long long f[4];
                    const int nl = min(min((int)(room - k), C - d), 4);
                    const Rec *grp = t.log + idx + k;           // k moves on below: fix the group's base
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        rh[j] = 0; m[j] = 0; rc[j] = 0; f[j] = 0; hh[j] = 0;
                        if (j < nl) {
                            hh[j] = hs[d + j];
                            ld_rec(grp + j, rh[j], m[j], f[j], rc[j]);
                        }
                    }
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        if (j < nl && !left) {
                            while (!rec_whole(rh[j], rc[j])) ld_rec_cg(grp + j, rh[j], m[j], f[j], rc[j]);
                            if (rh[j] != hh[j]) {
                                left = true;                    // the path leaves the run here
                            } else {
                                if (m[j] & a.avail) {
                                    if (allpre) { npre++; sel = m[j] & a.avail; }
                                } else {
                                    allpre = false;
                                    sl[d] = idx + k;
                                    if (f[j] < mine) atomicMax(&t.log[idx + k].first, mine);
                                }
                                d++; k++;
                            }
                        }
                    }
                }
                pidx = idx + k - 1;
            }
            // ---- create what is missing: one run per request, log space per warp
            const uint32_t L = (todo && d < C) ? (uint32_t)(C - d) : 0u;
            uint32_t incl = L;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t v = __shfl_up_sync(FULLW, incl, o);
                if (lane >= o) incl += v;
            }
            const uint32_t total = __shfl_sync(FULLW, incl, 31);
            unsigned long long wb = 0;
            if (total && lane == 0) wb = atomicAdd(&t.dev->log_count, (unsigned long long)total);
            wb = __shfl_sync(FULLW, wb, 0);
            bool lost = false;
            if (total && wb + total > (unsigned long long)t.log_cap) {
                if (lane == 0) atomicOr(&t.dev->status, B200R_S_TABLE_FULL);
                for (int x = d; L && x < C; x++) sl[x] = SLOT_NONE;
            } else if (L) {
                const uint32_t nb = (uint32_t)wb + incl - L;
                uint64_t oa = 0, ob = 1;
                if (eslot != SLOT_NONE)         // enter the run at the empty slot the probe ended at
                    cas128(t.tab + eslot, hd, ((uint64_t)parent_key(pidx) << 32) | nb, oa, ob);
                for (uint32_t k0 = 0; k0 < L; k0 += 4) {        // chunk hashes four at a time
                    uint64_t hh[4];
#pragma unroll
                    for (int j = 0; j < 4; j++) hh[j] = (k0 + j < L) ? hs[d + k0 + j] : 0;
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const uint32_t k = k0 + j;
                        if (k < L) {
                            st_rec_key(t.log + nb + k, hh[j], L - k);
                            atomicMax(&t.log[nb + k].first, mine);
                            sl[d + k] = nb + k;
                        }
                    }
                }
                lost = (oa | ob) != 0;          // the slot was taken meanwhile (rare): walk on from
            }                                   // depth d with coherent probes; the records stay unused
            if (!__any_sync(FULLW, lost)) break;
            todo = lost;
            coherent = true;
        }
        if (active) {
            a.dpre[i] = npre;
            a.cand[i] = sel;
        }
    }
}

}  // namespace b200r
