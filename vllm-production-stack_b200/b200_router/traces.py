"""Synthetic request traces for the BASELINE.json configs.

Deterministic (seeded numpy / random.Random); used by tests/ and bench.py.
The prompt shapes mirror the reference's benchmark client
(benchmarks/multi-round-qa/multi-round-qa.py:214-233,248-249: system prompt,
per-user context, numbered questions; src/tests/perftest/fake-openai-server.py:75
for the "Hello " answers) and the router-side flattening of chat messages
(src/vllm_router/routers/routing_logic.py:389-407, newline join).
"""
import random

import numpy as np

ALIGN = 16  # packed layout: every request starts on a 16-byte boundary


def endpoint_urls(n: int, port: int = 8000):
    """service_discovery.py:651 formats pod endpoints as http://<ip>:<port>."""
    return ["http://10.0.%d.%d:%d" % (i // 250, i % 250 + 1, port) for i in range(n)]


# --------------------------------------------------------------------- C1 / C5
def multi_round_qa(users=1000, rounds=10, sys_len=1000, hist_len=200, answer_len=100):
    """Arrival-ordered flattened prompts of a multi-round QA run.

    Returns (prompts, user_ids).  Round r of user u carries the whole chat
    history (user / assistant turns joined by newline) like the reference
    router sees it; users take turns round-robin inside a round."""
    sys_text = " ".join(["hi"] * sys_len)
    ctx_text = " ".join(["hi"] * hist_len)
    answer = "Hello " * answer_len
    history = []
    for u in range(users):
        history.append("Hi, here's some system prompt: %s.For user %d, here are some other "
                       "context: %s." % (sys_text, u, ctx_text))
    prompts, uids = [], []
    for r in range(rounds):
        q = ("Here's question #%d: can you tell me a new long story with a happy ending?"
             % (r + 1))
        for u in range(users):
            if r == 0:
                history[u] = history[u] + q
            else:
                history[u] = history[u] + "\n" + answer + "\n" + q
            prompts.append(history[u])
            uids.append(u)
    return prompts, uids


# -------------------------------------------------------------------------- C2
def session_ids_uuid(n: int, seed: int = 0):
    """uuid4-format lowercase ids (tests/e2e/test-routing.py:202 uses uuid4)."""
    rng = random.Random(seed)
    out = []
    for _ in range(n):
        x = rng.getrandbits(128)
        x = (x & ~(0xF << 76)) | (4 << 76)
        x = (x & ~(0x3 << 62)) | (2 << 62)
        h = "%032x" % x
        out.append("%s-%s-%s-%s-%s" % (h[:8], h[8:12], h[12:16], h[16:20], h[20:]))
    return out


def session_ids_packed(n: int, seed: int = 0):
    """Vectorised variant for the 1M-key config: 36-byte uuid-shaped ascii ids,
    packed at a 48-byte pitch.  Returns (bytes u8, starts i64, lens i32)."""
    rng = np.random.default_rng(seed)
    hexd = np.frombuffer(b"0123456789abcdef", dtype=np.uint8)
    body = hexd[rng.integers(0, 16, size=(n, 36), dtype=np.uint8)]
    body[:, [8, 13, 18, 23]] = ord("-")
    body[:, 14] = ord("4")
    body[:, 19] = hexd[8 + rng.integers(0, 4, size=n)]
    buf = np.zeros((n, 48), dtype=np.uint8)
    buf[:, :36] = body
    starts = np.arange(n, dtype=np.int64) * 48
    lens = np.full(n, 36, dtype=np.int32)
    return buf.reshape(-1), starts, lens


# -------------------------------------------------------------------------- C3
class SharedPrefixTrace:
    """N prompts = one of `n_sys` fixed system prompts (sys_chars ASCII chars)
    followed by a unique suffix of suffix_lo..suffix_hi chars over a 27-symbol
    alphabet; served in batches.  System prompt drawn uniformly."""

    def __init__(self, n_total=1 << 20, batch=1 << 16, n_sys=64, sys_chars=2048,
                 suffix_lo=256, suffix_hi=1024, seed=0):
        self.n_total, self.batch = n_total, batch
        self.n_sys, self.sys_chars = n_sys, sys_chars
        self.suffix_lo, self.suffix_hi = suffix_lo, suffix_hi
        self.seed = seed
        rng = np.random.default_rng(seed)
        alpha = np.frombuffer(b"abcdefghijklmnopqrstuvwxyz ", dtype=np.uint8)
        self.alpha = alpha
        self.sys = alpha[rng.integers(0, 27, size=(n_sys, sys_chars))]

    def num_batches(self):
        return (self.n_total + self.batch - 1) // self.batch

    def batch_arrays(self, b: int):
        """(bytes u8, starts i64, lens i32) of batch b; starts 16-byte aligned."""
        lo = b * self.batch
        n = min(self.batch, self.n_total - lo)
        rng = np.random.default_rng([self.seed, 1 + b])
        which = rng.integers(0, self.n_sys, size=n)
        slen = rng.integers(self.suffix_lo, self.suffix_hi + 1, size=n)
        lens = (self.sys_chars + slen).astype(np.int32)
        pitch = (lens.astype(np.int64) + ALIGN - 1) // ALIGN * ALIGN
        starts = np.zeros(n, dtype=np.int64)
        starts[1:] = np.cumsum(pitch[:-1])
        total = int(starts[-1] + pitch[-1]) if n else 0
        buf = np.zeros(total, dtype=np.uint8)
        pool = self.alpha[rng.integers(0, 27, size=int(slen.sum()))]
        sc = self.sys_chars
        po = 0
        for i in range(n):
            s = int(starts[i])
            buf[s:s + sc] = self.sys[which[i]]
            l = int(slen[i])
            buf[s + sc:s + sc + l] = pool[po:po + l]
            po += l
        return buf, starts, lens


# -------------------------------------------------------------------------- C4
def long_context_batch(n=8192, chars=32768, group=64, shared_frac=0.5, seed=0, chunk=128):
    """Prompts of exactly `chars` chars; members of a group of `group` prompts
    share the leading shared_frac of their chunks."""
    rng = np.random.default_rng(seed)
    alpha = np.frombuffer(b"abcdefghijklmnopqrstuvwxyz ", dtype=np.uint8)
    shared = int(round(chars // chunk * shared_frac)) * chunk
    buf = alpha[rng.integers(0, 27, size=(n, chars))]
    ngroups = (n + group - 1) // group
    heads = alpha[rng.integers(0, 27, size=(ngroups, shared))]
    if shared:
        buf[:, :shared] = heads[np.arange(n) // group]
    starts = np.arange(n, dtype=np.int64) * chars
    lens = np.full(n, chars, dtype=np.int32)
    return buf.reshape(-1), starts, lens


# ---------------------------------------------------------------- small / tests
def pack_strings(prompts, align: int = ALIGN):
    """Pack python strings: UTF-8 bytes, aligned starts, byte lens, char counts."""
    blobs = [p.encode("utf-8") for p in prompts]
    n = len(blobs)
    lens = np.array([len(b) for b in blobs], dtype=np.int32)
    nchars = np.array([len(p) for p in prompts], dtype=np.int32)
    pitch = (lens.astype(np.int64) + align - 1) // align * align
    starts = np.zeros(n, dtype=np.int64)
    if n > 1:
        starts[1:] = np.cumsum(pitch[:-1])
    total = int(starts[-1] + pitch[-1]) if n else 0
    buf = np.zeros(max(total, align), dtype=np.uint8)
    for i, b in enumerate(blobs):
        s = int(starts[i])
        buf[s:s + len(b)] = np.frombuffer(b, dtype=np.uint8)
    return buf, starts, lens, nchars


def sharing_stress(n_batches=6, batch=40, n_endpoints=6, max_depth=5, alphabet=3, seed=0,
                   chunk=128):
    """Heavy-sharing traces with availability churn: prompts
    are sequences of <= max_depth blocks drawn from a tiny alphabet, so most
    requests of a batch collide on trie nodes; each batch has its own
    availability subset.  Returns [(avail_indices, [prompts])]."""
    rng = random.Random(seed)
    out = []
    for _ in range(n_batches):
        k = rng.randint(1, n_endpoints)
        avail = sorted(rng.sample(range(n_endpoints), k))
        prompts = []
        for _ in range(batch):
            d = rng.randint(0, max_depth)
            s = "".join(chr(ord("a") + rng.randrange(alphabet)) * chunk for _ in range(d))
            tail = rng.choice([0, 0, 1, 17, chunk - 1])
            s += "z" * tail
            prompts.append(s)
        out.append((avail, prompts))
    return out
