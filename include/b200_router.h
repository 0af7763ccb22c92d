/*
 * b200_router.h -- C ABI of libb200router.so (hand-written sm_100a CUDA).
 *
 * Drop-in boundary for the routing decision of vllm_router (reference paths are
 * relative to the reference repository root):
 *
 *   prefix-aware routing   src/vllm_router/prefix/hashtrie.py:24-103
 *                          src/vllm_router/routers/routing_logic.py:347-423
 *   session routing        src/vllm_router/routers/routing_logic.py:84-103,170-219
 *                          (+ third-party uhashring==2.3 HashRing it calls)
 *
 * Conventions
 *   - Every pointer named d_* is DEVICE memory owned by the caller (e.g. a torch
 *     tensor's data_ptr()).  Nothing here takes or returns torch types.
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream).  The
 *     batch entry points only enqueue work: no host synchronisation, no
 *     allocation.  Object create/destroy/dump/status calls do synchronise.
 *   - Return value: 0 on success, a negative B200R_E_* code otherwise;
 *     b200r_last_error() returns a thread-local message for the last failure.
 *   - Errors detected on the device (table full, chunk capacity exceeded) are
 *     latched in the object and reported by b200r_trie_status().
 *   - There is no CPU fallback: without a CUDA device every compute entry point
 *     fails with B200R_E_CUDA.
 *
 * Packed prompt layout ("prompt bytes plus offsets"): request i occupies
 * d_bytes[d_starts[i] .. d_starts[i] + d_lens[i]) as UTF-8.  Requests whose
 * start address is 16-byte aligned and that are pure ASCII take the fast path
 * (cp.async staging, 16-byte vector loads); anything else is still hashed
 * bit-exactly by a slower generic path.  d_nchars[i] is the number of Unicode
 * code points of request i (Python's len(str)); chunking is by code points,
 * hashing over the UTF-8 bytes, exactly like hashtrie.py:55-56.
 */
#ifndef B200_ROUTER_H
#define B200_ROUTER_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default) /* the library is built with -fvisibility=hidden */
#endif

#define B200R_ABI_VERSION 2

#define B200R_OK 0
#define B200R_E_INVALID (-1)  /* bad argument                                  */
#define B200R_E_CUDA (-2)     /* CUDA runtime error (message has the detail)    */
#define B200R_E_NOMEM (-3)    /* device allocation failed                       */
#define B200R_E_SCRATCH (-4)  /* scratch buffer too small                       */
#define B200R_E_CAPACITY (-5) /* latched: table full / chunk capacity exceeded  */

/* flags for the prompt entry points */
#define B200R_F_ALL_ASCII 1u /* caller guarantees nchars[i] == lens[i] for all i */
/* b200r_prefix_hash_stage only: split it once more, so that three batches can be in flight.
 * PREP_ONLY runs everything but the chunk-hashing kernel (chunk counts, CSR, tile map, UTF-8
 * chunk offsets, ragged last chunks: latency-bound, a few MB); PREP_DONE runs only the
 * chunk-hashing kernel on a scratch buffer that a PREP_ONLY call has prepared. */
#define B200R_F_PREP_ONLY 2u
#define B200R_F_PREP_DONE 4u

/* latched device status bits (b200r_trie_status) */
#define B200R_S_TABLE_FULL 1u
#define B200R_S_CHUNK_CAP 2u
#define B200R_S_PEER_TIMEOUT 4u /* multi-GPU: a peer's data did not arrive within ~3 s */
#define B200R_S_INTERNAL 8u     /* internal: a run head was published without its record     */

typedef struct b200r_trie b200r_trie; /* GPU-resident HashTrie replacement     */
typedef struct b200r_ring b200r_ring; /* GPU-resident consistent-hash ring     */

int b200r_abi_version(void);
const char *b200r_last_error(void);

/* ------------------------------------------------------------------------
 * a1  HashTrie._chunk_and_hash                      hashtrie.py:45-56
 * For every request: split into chunks of `chunk_chars` code points, XXH64
 * (seed 0) each chunk.  Output is CSR: d_chunk_base[n+1] (int32, exclusive
 * prefix sum of per-request chunk counts) and d_hashes[d_chunk_base[n]].
 * max_chunks = capacity of d_hashes (host upper bound: sum ceil(len_i/chunk)).
 * d_nchars may be NULL: the library then counts code points on the device.
 * ---------------------------------------------------------------------- */
size_t b200r_xxh64_chunks_scratch_bytes(int64_t n, int64_t max_chunks);
int b200r_xxh64_chunks(const uint8_t *d_bytes, const int64_t *d_starts, const int32_t *d_lens,
                       const int32_t *d_nchars, int64_t n, int32_t chunk_chars, uint32_t flags,
                       int32_t *d_chunk_base, uint64_t *d_hashes, int64_t max_chunks,
                       void *d_scratch, size_t scratch_bytes, void *stream);

/* ------------------------------------------------------------------------
 * b200r_xxh64_chunks_noalloc -- the same chunk hashing for callers that cannot
 * allocate.  (No counterpart in the reference; added for embedders.)
 *
 * Who needs it: code that is only ever given input and output buffers -- a
 * plugin inside a serving process whose engine has already claimed almost all
 * device memory (a cudaMalloc for work space would fail), or a request-path
 * thread that must avoid cudaMalloc/cudaFree because they stall the device.
 * The work space is a fixed B200R_NOALLOC_SCRATCH_BYTES area compiled into
 * the library (each GPU that loads the library holds it until exit); the call
 * itself performs no cudaMalloc/cudaFree and uses no allocating routine.
 *
 * Same arguments and results as b200r_xxh64_chunks minus the scratch pair.
 * Limits and rules:
 *   - the batch must fit the built-in area: roughly 4*n + 4*max_chunks bytes
 *     for non-ASCII batches (about 900k chunks = 115 MB of prompt text), far
 *     more with B200R_F_ALL_ASCII; larger batches fail with B200R_E_SCRATCH
 *     and a message giving the required size;
 *   - the first call on each device looks up the area's address (with that
 *     device current); make that first call before starting any stream capture;
 *   - all calls on a device share the one area, which keeps whatever the last
 *     call left in it (nothing needs to be cleared between calls); two calls
 *     whose kernels may overlap on the GPU -- different streams without an
 *     ordering between them -- race on it and must be serialised by the caller.
 * ---------------------------------------------------------------------- */
#define B200R_NOALLOC_SCRATCH_BYTES (4 * 1024 * 1024)
int b200r_xxh64_chunks_noalloc(const uint8_t *d_bytes, const int64_t *d_starts,
                               const int32_t *d_lens, const int32_t *d_nchars, int64_t n,
                               int32_t chunk_chars, uint32_t flags, int32_t *d_chunk_base,
                               uint64_t *d_hashes, int64_t max_chunks, void *stream);

/* ------------------------------------------------------------------------
 * a4/a6  TrieNode / HashTrie.__init__               hashtrie.py:24-43
 * The pointer trie becomes a node log plus a hash table of run heads.  Every
 * node is one 32-byte record {chunk hash, endpoint bitmask, batch bookkeeping}
 * and is identified by its index in the log.  The nodes one request creates in
 * one go are consecutive records (a run); only the first node of a run is
 * entered in an open-addressing table keyed by (parent node index, chunk hash)
 * -- exactly parent.children[chunk_hash] -- with the log index as value.
 * Walking a stored path compares consecutive records with the request's chunk
 * hashes.  Capacity = 2^cap_log2 nodes (log records of 32 bytes) and
 * 2^(cap_log2+1) table entries of 16 bytes (load factor <= 0.5 even if every
 * node heads its own run): 64 bytes of device memory per node of capacity.  Endpoints are bit
 * indices 0..63.  The reference never evicts, neither does the log:
 * b200r_trie_status reports the node count; a full log latches
 * B200R_S_TABLE_FULL.
 * ---------------------------------------------------------------------- */
int b200r_trie_create(int32_t cap_log2, int32_t chunk_chars, b200r_trie **out);
int b200r_trie_destroy(b200r_trie *t);
int b200r_trie_clear(b200r_trie *t, void *stream);
/* synchronises `stream`; any of the outputs may be NULL */
int b200r_trie_status(b200r_trie *t, void *stream, uint32_t *status_bits, int64_t *num_nodes,
                      uint64_t *root_mask, int64_t *capacity);
/* The latched status word copied to a DEVICE word in stream order (no host sync), so that a
 * caller can bring it back in the same device-to-host copy as the batch's decisions. */
int b200r_trie_status_async(b200r_trie *t, uint32_t *d_status_out, void *stream);
/* a11 endpoint indexing (service_discovery.py:651 mints a new url per pod restart; the
 * reference keeps stale urls in its sets forever and filters them by intersection with the
 * available set, hashtrie.py:96).  Masks are one 64-bit word, so the bit of an endpoint that
 * is gone for good has to be recycled: this sweep clears `bits` from every endpoint set and
 * from the root set.  Decisions are unchanged as long as those endpoints never become
 * available again under the same bit. */
int b200r_trie_retire_endpoints(b200r_trie *t, uint64_t bits, void *stream);
/* Test hook: set the batch epoch (the counter behind the per-node "first toucher" words).
 * The epoch advances once per routed batch; when its 32 bits run out the library zeroes
 * those words with one sweep and restarts at 1 -- nothing else changes. */
int b200r_trie_set_epoch(b200r_trie *t, uint32_t epoch);

/* ------------------------------------------------------------------------
 * a2  HashTrie.longest_prefix_match                 hashtrie.py:76-103
 * Frozen-table lookup from chunk hashes (CSR as produced by a1).  Outputs per
 * request: d_match_depth = matched chunks (reference match_length / chunk_size),
 * d_matched_mask = selected endpoint set (avail_mask itself when depth is 0).
 * ---------------------------------------------------------------------- */
int b200r_trie_lookup_batch(b200r_trie *t, const uint64_t *d_hashes, const int32_t *d_chunk_base,
                            int64_t n, uint64_t avail_mask, int32_t *d_match_depth,
                            uint64_t *d_matched_mask, void *stream);

/* ------------------------------------------------------------------------
 * a3  HashTrie.insert                               hashtrie.py:58-74
 * Adds d_endpoint[i] to the root and to every node on request i's path,
 * creating missing nodes.  Order-independent (set union), so equal to the
 * reference's sequential inserts.  Entries with d_endpoint[i] < 0 are skipped.
 * ---------------------------------------------------------------------- */
int b200r_trie_insert_batch(b200r_trie *t, const uint64_t *d_hashes, const int32_t *d_chunk_base,
                            int64_t n, const int32_t *d_endpoint, void *stream);

/* ------------------------------------------------------------------------
 * a5  PrefixAwareRouter.route_request               routing_logic.py:414-423
 * Fused batch: hash, lookup, tie-break, insert, with results equal to routing
 * the n requests one after another in index order against the same table
 * (SURVEY.md 7.3 H1).  avail_mask must be constant for the batch.  Tie-break:
 * candidates in endpoint-index order, draw = b200r_pick(seed, seq0 + i, count).
 * Outputs per request: d_endpoint (chosen bit index), d_match_depth,
 * d_matched_mask (the candidate set the request drew from).
 * ---------------------------------------------------------------------- */
size_t b200r_prefix_route_scratch_bytes(int64_t n, int64_t max_chunks, uint32_t flags);
int b200r_prefix_route_batch(b200r_trie *t, const uint8_t *d_bytes, const int64_t *d_starts,
                             const int32_t *d_lens, const int32_t *d_nchars, int64_t n,
                             int64_t max_chunks, uint32_t flags, uint64_t avail_mask,
                             uint64_t seed, uint64_t seq0, int32_t *d_endpoint,
                             int32_t *d_match_depth, uint64_t *d_matched_mask, void *d_scratch,
                             size_t scratch_bytes, void *stream);
/* The same batch as two stages that meet only in the scratch buffer, for callers
 * that overlap batches: the hash stage (prep + chunk hashing; never touches the
 * table) of batch b+1 can run on another stream while the route stage (claim /
 * select / commit) of batch b runs; one scratch buffer per batch in flight.
 * Route stages must be issued in batch order.  b200r_prefix_route_batch is
 * exactly hash stage + route stage on one stream. */
int b200r_prefix_hash_stage(b200r_trie *t, const uint8_t *d_bytes, const int64_t *d_starts,
                            const int32_t *d_lens, const int32_t *d_nchars, int64_t n,
                            int64_t max_chunks, uint32_t flags, void *d_scratch,
                            size_t scratch_bytes, void *stream);
int b200r_prefix_route_stage(b200r_trie *t, int64_t n, int64_t max_chunks, uint32_t flags,
                             uint64_t avail_mask, uint64_t seed, uint64_t seq0,
                             int32_t *d_endpoint, int32_t *d_match_depth,
                             uint64_t *d_matched_mask, void *d_scratch, size_t scratch_bytes,
                             void *stream);
/* The same three steps -- prep | hash | route -- pipelined over consecutive batches by the
 * library itself, on three streams the pipeline owns (prep and route at the highest stream
 * priority, the SM-filling hashing kernel at the lowest): prep of batch b+2 and hash of batch
 * b+1 run while route of batch b runs; route stages stay in call order, so results are those
 * of sequential routing, and the caller's `stream` waits for the call's results.  The caller rotates `depth` scratch buffers
 * (call i uses buffer i mod depth; the pipeline makes call i+depth wait for call i).
 * inputs_on_prep_stream = 0: the inputs were produced on `stream` (the prep step waits for
 * it); 1: they are complete, or were copied on b200r_pipe_prep_stream(), e.g. a host-to-device
 * copy that should itself overlap the routing of earlier batches.  (The reference routes one
 * request at a time, request.py:252-257; there is no counterpart.) */
typedef struct b200r_pipe b200r_pipe;
int b200r_pipe_create(b200r_trie *t, int32_t depth, b200r_pipe **out);
int b200r_pipe_destroy(b200r_pipe *p);
void *b200r_pipe_prep_stream(b200r_pipe *p);
/* Optional, before copying the next call's inputs on the prep stream into memory that an
 * earlier call's inputs occupied: makes the prep stream wait until the call that used this
 * rotation slot `depth` calls ago has finished (it no longer reads its inputs then). */
int b200r_pipe_begin(b200r_pipe *p);
/* host-to-device copy of inputs on the prep stream (pinned host memory for a true async copy) */
int b200r_pipe_copy_h2d(b200r_pipe *p, void *d_dst, const void *h_src, size_t bytes);
int b200r_pipe_route(b200r_pipe *p, const uint8_t *d_bytes, const int64_t *d_starts,
                     const int32_t *d_lens, const int32_t *d_nchars, int64_t n, int64_t max_chunks,
                     uint32_t flags, uint64_t avail_mask, uint64_t seed, uint64_t seq0,
                     int32_t *d_endpoint, int32_t *d_match_depth, uint64_t *d_matched_mask,
                     void *d_scratch, size_t scratch_bytes, int32_t inputs_on_prep_stream, void *stream);
/* The route stage from chunk hashes the caller already holds (CSR exactly as b200r_xxh64_chunks
 * writes it): claim / select / commit, same results as b200r_prefix_route_batch on the prompts
 * those hashes came from.  Used where hashes of several sources are merged before routing
 * (replicated multi-GPU tables: every rank all-gathers the ranks' hashes and routes the global
 * batch).  Scratch: b200r_prefix_route_scratch_bytes(n, max_chunks, B200R_F_ALL_ASCII). */
int b200r_prefix_route_hashes(b200r_trie *t, const uint64_t *d_hashes, const int32_t *d_chunk_base,
                              int64_t n, int64_t max_chunks, uint64_t avail_mask, uint64_t seed,
                              uint64_t seq0, int32_t *d_endpoint, int32_t *d_match_depth,
                              uint64_t *d_matched_mask, void *d_scratch, size_t scratch_bytes,
                              void *stream);
/* the canonical draw, on the host (same arithmetic as the kernels) */
uint32_t b200r_pick(uint64_t seed, uint64_t seq, uint32_t n);

/* ------------------------------------------------------------------------
 * Trie state export (parity / checkpoint; the reference has no counterpart
 * beyond walking TrieNode.children, hashtrie.py:24-32).  Writes up to
 * max_nodes entries of 4 x u64
 *   {log index | records left in its run << 32,
 *    parent's log index (0xFFFFFFFF = the root) | 1 << 63 on the first node of a run,
 *    chunk hash, mask}
 * to d_out and the number of nodes to *count (may exceed max_nodes).
 * Synchronises.  Paths follow from the parent links.
 * ---------------------------------------------------------------------- */
int b200r_trie_dump(b200r_trie *t, uint64_t *d_out, int64_t max_nodes, int64_t *count,
                    uint64_t *root_mask, void *stream);
/* Checkpoint restore: writes dumped entries back at their log indices and their run
 * heads back into the table, so the restored table has the layout of the saved one.
 * Node identity is the log position, hence only into an EMPTY table (fresh or cleared)
 * of at least the dump's capacity; otherwise B200R_E_INVALID.  Synchronises once. */
int b200r_trie_restore(b200r_trie *t, const uint64_t *d_entries, int64_t count,
                       uint64_t root_mask, void *stream);

/* Latency path: a5 for ONE request with HOST buffers in and out -- one launch (the prompt is
 * read from a pinned staging buffer of the table straight into shared memory, hashed, walked,
 * the endpoint drawn and inserted) and one stream synchronisation; same results as a batch of
 * one.  Returns 0 = routed; 1 = not taken (prompt not pure ASCII, longer than 16384 chunks, or
 * a chunk size that is not a multiple of 16): route it as a batch of one instead.  `status`
 * receives the latched B200R_S_* bits.  The staging buffer and ~200 KB of device scratch are
 * allocated on the table's first call (not a no-allocation entry point).
 * Replaces PrefixAwareRouter.route_request for a lone request (routing_logic.py:363-423). */
int b200r_prefix_route_one(b200r_trie *t, const uint8_t *h_prompt, int64_t len_bytes,
                           uint64_t avail_mask, uint64_t seed, uint64_t seq, int32_t *endpoint,
                           int32_t *match_depth, uint64_t *matched_mask, uint32_t *status,
                           void *stream);

/* ------------------------------------------------------------------------
 * Multi-GPU with a PARTITIONED table (scalable variant; no counterpart in the
 * reference).  A request's whole path hangs under its first chunk, so the
 * table is split by owner = f(first chunk hash) % world and nothing is
 * replicated: every rank hashes its own shard, ships the chunk hashes (8 bytes
 * per 128-byte chunk) of each request to the rank owning it, routes the
 * requests it owns in global order against its partition and ships the results
 * back -- all through NVLink peer memory by this library's kernels, no host
 * round trip.  Results equal sequential routing of the rank-major global batch.
 * Buffers: one peer-addressable buffer of b200r_part_layout(...)[0] bytes per
 * rank (zeroed once, before the first batch) and a private work buffer of
 * [1] bytes (zeroed once); d_peer_bases = u64[world] base addresses.  cap /
 * rcap = per-rank chunk / request capacity of a batch (both even: the exchange
 * ships 16-byte vectors); epoch = 1, 2, 3, ...
 * Per batch and rank: b200r_part_send, b200r_part_route, b200r_part_finish.
 * ---------------------------------------------------------------------- */
int b200r_part_layout(int32_t world, int64_t cap, int64_t rcap, size_t out[2]);
int b200r_part_send(b200r_trie *t, const uint8_t *d_bytes, const int64_t *d_starts,
                    const int32_t *d_lens, const int32_t *d_nchars, int64_t n, int64_t max_chunks,
                    uint32_t flags, const uint64_t *d_peer_bases, int32_t world, int32_t rank,
                    int64_t cap, int64_t rcap, uint32_t epoch, void *d_work, void *d_scratch,
                    size_t scratch_bytes, void *stream);
int b200r_part_route(b200r_trie *t, uint64_t avail_mask, uint64_t seed,
                     const uint64_t *d_peer_bases, void *d_my_base, int32_t world, int32_t rank,
                     int64_t cap, int64_t rcap, uint32_t epoch, void *d_work, void *stream);
int b200r_part_finish(b200r_trie *t, int64_t n, void *d_my_base, int32_t world, int32_t rank,
                      int64_t cap, int64_t rcap, uint32_t epoch, void *d_work, int32_t *d_endpoint,
                      int32_t *d_match_depth, uint64_t *d_matched, void *stream);

/* The three phases of one batch in ONE call, pipelined over consecutive batches on four
 * streams the table owns: prep (input copies, per-request counting + scan), send (the
 * SM-filling hashing kernel, placement and shipment; lowest stream priority), route (wait for
 * the sources, gather, claim, select, commit) and return (push results, wait for the owners,
 * scatter); prep, route and return run at the highest priority.  Send of batch
 * b+1 overlaps route and return of batch b; route of batch b+1 follows the commit of batch b.
 * The caller's `stream` waits for the call's results.  d_scratch must be one of THREE buffers
 * used in turn (batch b: buffer b % 3, like the exchange's buffer sets; the prep of batch b+1
 * runs while batch b is being hashed).
 * inputs_on_prep_stream = 1: the inputs are complete or were copied with
 * b200r_part_copy_h2d after b200r_part_begin(t, epoch) (a host-to-device copy that overlaps the
 * previous batch); 0: they were produced on `stream`. */
int b200r_part_begin(b200r_trie *t, uint32_t epoch);
int b200r_part_copy_h2d(b200r_trie *t, void *d_dst, const void *h_src, size_t bytes);
int b200r_part_step(b200r_trie *t, const uint8_t *d_bytes, const int64_t *d_starts,
                    const int32_t *d_lens, const int32_t *d_nchars, int64_t n, int64_t max_chunks,
                    uint32_t flags, uint64_t avail_mask, uint64_t seed, const uint64_t *d_peer_bases,
                    void *d_my_base, int32_t world, int32_t rank, int64_t cap, int64_t rcap,
                    uint32_t epoch, void *d_work, void *d_scratch, size_t scratch_bytes,
                    int32_t *d_endpoint, int32_t *d_match_depth, uint64_t *d_matched,
                    int32_t inputs_on_prep_stream, void *stream);

/* ------------------------------------------------------------------------
 * a9  uhashring.HashRing (md5 points, bisect_right + wrap), a7 SessionRouter
 *                                                   routing_logic.py:84-103,217
 * b200r_md5_batch: 128-bit digests as the big-endian integer uhashring uses,
 * d_out[2*i] = high 64 bits, d_out[2*i+1] = low 64 bits.
 * b200r_ring_build: upload `npoints` ring points sorted ascending with their
 * owner endpoint indices (HOST pointers; the ring is small).
 * b200r_ring_lookup_batch: owner of the first point strictly greater than
 * md5(key), wrapping to point 0; -1 when the ring is empty.
 * ---------------------------------------------------------------------- */
int b200r_md5_batch(const uint8_t *d_bytes, const int64_t *d_starts, const int32_t *d_lens,
                    int64_t n, uint64_t *d_out, void *stream);
int b200r_ring_create(b200r_ring **out);
int b200r_ring_destroy(b200r_ring *r);
int b200r_ring_build(b200r_ring *r, const uint64_t *h_points_hi, const uint64_t *h_points_lo,
                     const int32_t *h_owner, int32_t npoints, void *stream);
int b200r_ring_lookup_batch(b200r_ring *r, const uint8_t *d_bytes, const int64_t *d_starts,
                            const int32_t *d_lens, int64_t n, int32_t *d_endpoint, void *stream);

/* ------------------------------------------------------------------------
 * Measurement hooks (bench.py; no counterpart in the reference).
 * b200r_prof_enable(1): every pipeline kernel is bracketed by CUDA events on
 * its launch stream.  b200r_prof_read synchronises those events and returns,
 * per kernel class, the summed milliseconds and the number of launches since
 * the last read; slots: 0 prep, 1 utf8, 2 hash, 3 claim, 4 select, 5 commit,
 * 6 lookup, 7 insert, 8 classify, 9 push, 10 wait, 11 apply, 12 resolve.
 * b200r_launch_count: kernels enqueued by this library
 * since load (process-wide).
 * ---------------------------------------------------------------------- */
/* tuning switch of the route stage's claim step, i.e. how many lanes walk one request: 0 =
 * choose per batch from the average number of chunks (one thread up to 64 chunks per prompt, a
 * warp beyond), 1 = one thread, 2 = a warp, 3 = 16 lanes, 4 = 8 lanes; identical results */
int b200r_set_claim_mode(int mode);

/* Partitioned multi-GPU send phase: 0 (default) = for ASCII batches the hashing kernel stores the
 * chunk hashes straight into the owners' buffers (placement done by the prep kernel); 1 = always
 * hash into local CSR order, compact per owner and ship (the path non-ASCII batches take).  Same
 * results; for tests and comparison. */
int b200r_set_part_staged(int on);

/* Chunk-hashing kernel: 0 = persistent warps that stage tile t+2 while hashing tile t
 * (k_hash_stream), 1 = one 32-chunk tile per warp (k_hash_chunks).  Same results. */
int b200r_set_hash_mode(int mode);
#define B200R_PROF_SLOTS 13
int b200r_prof_enable(int on);
int b200r_prof_read(float *ms_out, long long *count_out, int nslots);
long long b200r_launch_count(void);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* B200_ROUTER_H */
