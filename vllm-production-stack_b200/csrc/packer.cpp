// Host-side prompt packer: Python strings -> the packed layout the
// routing kernels read (include/b200_router.h: bytes / starts / lens / nchars, every start
// 16-byte aligned), written straight into caller-owned (pinned) memory.
//
// Replaces the per-request work the reference does in Python before it can hash anything
// (routing_logic.py:388-412 builds the prompt string; hashtrie.py:45-56 slices and encodes it).
// A prompt is either one str or a sequence of str parts that the reference would have joined
// with a separator ("\n".join(contents), routing_logic.py:407): the parts are laid end to end
// with the separator between them, so the joined string is never materialised.
//
// This is the one file of the library that talks to CPython (it has to read str objects);
// it is built into its own shared object (libb200pack.so, no CUDA) and is called with the GIL
// held (ctypes.PyDLL).  Entry points return 0 or a negative error; b200p_last_error() says why.
#define PY_SSIZE_T_CLEAN
#include <Python.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

namespace {

thread_local char g_err[256] = "";

int fail(const char *msg)
{
    snprintf(g_err, sizeof g_err, "%s", msg);
    PyErr_Clear();
    return -1;
}

constexpr int64_t ALIGN = 16;

struct Piece { const char *p; Py_ssize_t bytes, chars; };

// UTF-8 view of a str.  ASCII strings are stored as ASCII by CPython: no copy, no cache.
inline bool view(PyObject *s, Piece *out)
{
    if (!PyUnicode_Check(s)) return false;
    if (PyUnicode_IS_COMPACT_ASCII(s)) {
        out->p = (const char *)PyUnicode_1BYTE_DATA(s);
        out->bytes = out->chars = PyUnicode_GET_LENGTH(s);
        return true;
    }
    Py_ssize_t nb = 0;
    const char *p = PyUnicode_AsUTF8AndSize(s, &nb);      // cached on the object by CPython
    if (!p) return false;
    out->p = p; out->bytes = nb; out->chars = PyUnicode_GET_LENGTH(s);
    return true;
}

// visits the pieces of prompt `item` (a str, or a list/tuple of str joined by sep)
template <typename F>
inline bool for_pieces(PyObject *item, const Piece &sep, F &&f)
{
    Piece pc;
    if (PyUnicode_Check(item)) {
        if (!view(item, &pc)) return false;
        f(pc);
        return true;
    }
    if (!PyList_Check(item) && !PyTuple_Check(item)) return false;
    PyObject **parts = PySequence_Fast_ITEMS(item);
    const Py_ssize_t np = PySequence_Fast_GET_SIZE(item);
    for (Py_ssize_t k = 0; k < np; k++) {
        if (k) f(sep);
        if (!view(parts[k], &pc)) return false;
        f(pc);
    }
    return true;
}

}  // namespace

extern "C" {

__attribute__((visibility("default"))) const char *b200p_last_error(void) { return g_err; }

// Pass 1: sizes.  prompts = list or tuple; sep = str (used between the parts of multi-part
// prompts) or None.  out[0] = bytes of the packed buffer (with alignment padding, >= 16),
// out[1] = payload bytes, out[2] = sum over prompts of ceil(code points / chunk_chars),
// out[3] = 1 if every prompt is pure ASCII.  by_bytes != 0: chunks are counted in bytes.
__attribute__((visibility("default")))
int b200p_measure(PyObject *prompts, PyObject *sep_obj, int64_t chunk_chars, int by_bytes, int64_t out[4])
{
    if (!prompts || !(PyList_Check(prompts) || PyTuple_Check(prompts))) return fail("prompts must be a list or tuple");
    if (chunk_chars <= 0) return fail("chunk_chars must be positive");
    Piece sep = { "", 0, 0 };
    if (sep_obj && sep_obj != Py_None && !view(sep_obj, &sep)) return fail("separator must be a str");
    PyObject **items = PySequence_Fast_ITEMS(prompts);
    const Py_ssize_t n = PySequence_Fast_GET_SIZE(prompts);
    int64_t total = 0, payload = 0, chunks = 0;
    bool ascii = true;
    for (Py_ssize_t i = 0; i < n; i++) {
        int64_t nb = 0, nc = 0;
        if (!for_pieces(items[i], sep, [&](const Piece &pc) { nb += pc.bytes; nc += pc.chars; }))
            return fail("every prompt must be a str or a list/tuple of str");
        if (nb > 0x7fffffff) return fail("prompt longer than 2 GiB");
        payload += nb;
        total += (nb + ALIGN - 1) / ALIGN * ALIGN;
        chunks += ((by_bytes ? nb : nc) + chunk_chars - 1) / chunk_chars;
        ascii = ascii && nb == nc;
    }
    out[0] = total > 0 ? total : ALIGN;
    out[1] = payload;
    out[2] = chunks;
    out[3] = ascii ? 1 : 0;
    return 0;
}

// Pass 2: copy.  buf has out[0] bytes from b200p_measure; starts/lens/nchars have n entries.
// Padding bytes are zeroed.  by_bytes != 0: nchars[i] = lens[i] (chunk by bytes, the
// gateway picker's semantics, prefix_aware_picker.go:154-165).
// threads: worker threads for the copies (0 = pick: one per ~8 MB, at most 8; 1 = none).  The
// Python objects are only read by the calling thread, with the GIL held; the copies work on raw
// (pointer, length, destination) triples and run with the GIL RELEASED -- the triples stay
// valid as long as the caller keeps the list (and so the strings) alive and unmodified
// until the call returns.
__attribute__((visibility("default")))
int b200p_fill(PyObject *prompts, PyObject *sep_obj, uint8_t *buf, int64_t buf_bytes, int64_t *starts,
               int32_t *lens, int32_t *nchars, int by_bytes, int threads)
{
    if (!prompts || !(PyList_Check(prompts) || PyTuple_Check(prompts))) return fail("prompts must be a list or tuple");
    Piece sep = { "", 0, 0 };
    if (sep_obj && sep_obj != Py_None && !view(sep_obj, &sep)) return fail("separator must be a str");
    PyObject **items = PySequence_Fast_ITEMS(prompts);
    const Py_ssize_t n = PySequence_Fast_GET_SIZE(prompts);
    if (!buf || (n > 0 && (!starts || !lens || !nchars))) return fail("null output buffer");
    struct Job { const char *src; int64_t bytes, dst; };      // src == nullptr: zero fill
    std::vector<Job> jobs;
    jobs.reserve((size_t)n + (size_t)n / 2 + 4);
    int64_t off = 0;
    for (Py_ssize_t i = 0; i < n; i++) {
        int64_t nb = 0, nc = 0;
        const bool ok = for_pieces(items[i], sep, [&](const Piece &pc) {
            if (pc.bytes) jobs.push_back({ pc.p, (int64_t)pc.bytes, off + nb });
            nb += pc.bytes; nc += pc.chars;
        });
        if (!ok) return fail("every prompt must be a str or a list/tuple of str");
        const int64_t pitch = (nb + ALIGN - 1) / ALIGN * ALIGN;
        if (off + pitch > buf_bytes) return fail("packed buffer too small (prompts changed since b200p_measure?)");
        if (pitch > nb) jobs.push_back({ nullptr, pitch - nb, off + nb });
        starts[i] = off;
        lens[i] = (int32_t)nb;
        nchars[i] = (int32_t)(by_bytes ? nb : nc);
        off += pitch;
    }
    if (off == 0 && buf_bytes >= ALIGN) memset(buf, 0, ALIGN);
    auto run = [&](size_t a, size_t b) {
        for (size_t j = a; j < b; j++) {
            if (jobs[j].src) memcpy(buf + jobs[j].dst, jobs[j].src, (size_t)jobs[j].bytes);
            else memset(buf + jobs[j].dst, 0, (size_t)jobs[j].bytes);
        }
    };
    int T = threads;
    if (T <= 0) {
        T = (int)(off >> 23) + 1;
        const int hw = (int)std::thread::hardware_concurrency();
        if (T > 8) T = 8;
        if (hw > 0 && T > hw) T = hw;
    }
    // From here on no Python object is touched: let other Python threads run (a second
    // batch can be measured / gathered while this one is being copied).
    PyThreadState *ts = PyEval_SaveThread();
    if (T <= 1 || jobs.size() < 64) {
        run(0, jobs.size());
        PyEval_RestoreThread(ts);
        return 0;
    }
    // equal byte shares: jobs are in destination order, so job j ends at dst + bytes
    std::vector<std::thread> pool;
    size_t a = 0;
    for (int t = 0; t < T; t++) {
        size_t b = a;
        const int64_t until = off / T * (t + 1);
        if (t == T - 1) b = jobs.size();
        else while (b < jobs.size() && jobs[b].dst + jobs[b].bytes <= until) b++;
        if (b > a) pool.emplace_back(run, a, b);
        a = b;
    }
    for (auto &th : pool) th.join();
    PyEval_RestoreThread(ts);
    return 0;
}

}  // extern "C"
