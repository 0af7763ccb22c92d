/* b200_pack.h -- host-side prompt packer of the B200 router (libb200pack.so, no CUDA).
 *
 * Python strings -> the packed prompt layout of b200_router.h (bytes / starts / lens /
 * nchars, every start 16-byte aligned), written straight into caller-owned, normally
 * pinned, memory.  Stands where the reference builds the prompt string and slices /
 * encodes it per chunk (src/vllm_router/routers/routing_logic.py:388-412,
 * src/vllm_router/prefix/hashtrie.py:45-56).  Unlike b200_router.h this is a CPython
 * binding helper: the arguments are borrowed PyObject references and the calls must be
 * made with the GIL held (ctypes.PyDLL).  0 = ok, negative = error (b200p_last_error).
 */
#ifndef B200_PACK_H
#define B200_PACK_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
#pragma GCC visibility push(default)

struct _object; /* PyObject */

const char *b200p_last_error(void);

/* prompts: list/tuple whose items are str, or list/tuple of str parts that are joined by
 * `sep` (str or None) -- "\n".join(contents) of routing_logic.py:407 without building the
 * joined string.  out[0] = bytes of the packed buffer (>= 16), out[1] = payload bytes,
 * out[2] = sum of ceil(units / chunk_chars) (units = code points, or bytes if by_bytes),
 * out[3] = 1 if every prompt is pure ASCII. */
int b200p_measure(struct _object *prompts, struct _object *sep, int64_t chunk_chars, int by_bytes,
                  int64_t out[4]);

/* buf: out[0] bytes; starts/lens/nchars: one entry per prompt.  Padding is zeroed.
 * by_bytes != 0: nchars[i] = lens[i], i.e. chunks are 128 BYTES -- the chunking of the
 * gateway picker (src/gateway_inference_extension/prefix_aware_picker.go:154-165).
 * threads: worker threads for the copies (0 = automatic, at most 8; 1 = copy on the calling
 * thread); the Python objects themselves are only touched by the calling thread. */
int b200p_fill(struct _object *prompts, struct _object *sep, uint8_t *buf, int64_t buf_bytes,
               int64_t *starts, int32_t *lens, int32_t *nchars, int by_bytes, int threads);

#pragma GCC visibility pop
#ifdef __cplusplus
}
#endif
#endif
