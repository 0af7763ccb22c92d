#!/usr/bin/env python
"""bench.py -- prefix-route decisions/s (and prompt GB/s vs the HBM roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--workload prefix|session|mrqa|c4|latency]

Default workload (BASELINE.json configs[2], SURVEY.md 8d "C3"): prompts = one of 64 shared
2048-char system prompts + a unique 256..1024-char suffix, 8 endpoints, batches of 65536
requests per GPU, one endpoint dropped for two batches (availability churn).  A step = one
batch through the whole routing path (hash, lookup, tie-break, insert; results equal to the
reference's sequential order).  Every step routes a batch the table has not seen; batches
(~177 MB of prompt bytes) are larger than L2 (126 MB) and distinct, so no flush is needed
between steps.  The K timed steps are repeated over PASSES fresh tables (the same W warm-up +
K timed batches each) so that the timed total is tens of milliseconds; ms_per_step is the
mean over all timed steps, and per-pass values are listed.

  value      whole-job decisions/s, batches resident in HBM, CUDA events, max over ranks
  e2e        same through the public Python API from pinned HOST buffers:
             H2D of the packed batch + route + D2H of the decisions, per step
  roofline   the longest kernel of the step (algorithmic bytes of that kernel / its CUDA-event
             duration, single-stream pass with per-kernel events) plus whole_path: the
             SURVEY 8d bytes of a routed request, L + 8 + 32 C + 32 C_new + 4, over the step
  cpu_baseline  oracle/prefix_ref.py (Python restatement of the reference, asyncio
             structure included) on a bounded sample, 1 core
  parity_sample_ok  decisions / depths / candidate sets of churn-bearing batches AND the final
             trie equal the C oracle's (N>1: per rank against the rank-major global order,
             union of the partitions == oracle trie); no value is printed when it is false
  extras     C2 (session ring), C5 (multi-round-QA replay), C4 (long context, N>1: with
             exchange bytes / time) and single-request latency figures, each with clocks

--impl reference times the restatement alone (the reference is pure Python and is not
installable on the GPU box).  Under torchrun every rank runs the b200 arm on
its shard; for --impl reference only rank 0 works.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "vllm-production-stack_b200"))

import numpy as np  # noqa: E402

BATCH = 1 << 16
N_ENDPOINTS = 8
N_SYS = 64
SYS_CHARS = 2048
SUFFIX = (256, 1024)
SEED = 20240607
METRIC = "prefix_route_decisions_per_sec"
UNIT = "decisions/s"
PASSES = 5


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="prefix",
                    choices=["prefix", "session", "mrqa", "c4", "latency"],
                    help="prefix = BASELINE metric (default); session = configs[1], 1M session ids; "
                         "mrqa = configs[4], multi-round-QA trace replay; c4 = configs[3], 32K-char "
                         "prompts, 64 endpoints; latency = one request at a time")
    ap.add_argument("--batch", type=int, default=BATCH, help=argparse.SUPPRESS)
    ap.add_argument("--cpu-sample", type=int, default=BATCH, help=argparse.SUPPRESS)
    ap.add_argument("--no-cpu", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--no-extras", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--passes", type=int, default=PASSES, help=argparse.SUPPRESS)
    return ap.parse_args()


def workload_config(args, n_gpus):
    return {"workload": "C3: 64 shared 2048-char system prompts + unique 256-1024-char suffix, "
                        "%d endpoints, batch %d per GPU, endpoint 3 unavailable in batches 5-6"
                        % (N_ENDPOINTS, args.batch),
            "batch_per_gpu": args.batch, "global_batch": args.batch * n_gpus,
            "chunk_chars": 128, "endpoints": N_ENDPOINTS,
            "l2": "inputs larger than L2: each step reads a distinct ~%d MB batch"
                  % (args.batch * 2700 // 1000000),
            "passes": "%d passes of (warmup + steps) on fresh tables; ms_per_step = mean over all "
                      "timed steps" % args.passes,
            "tie_break": "canonical (endpoint-index order, counter-based draw)",
            "multi_gpu": None if n_gpus == 1 else
            {"": "table partitioned by first-chunk hash, NVLink peer-memory exchange",
             "replicated": "table replicated, NCCL all-gather of chunk hashes",
             "nccl": "table replicated, NCCL all-gather of chunk hashes"}[
                os.environ.get("BENCH_EXCHANGE", "")]}


def avail_mask(step_index):
    full = (1 << N_ENDPOINTS) - 1
    return full & ~(1 << 3) if step_index in (5, 6) else full


def make_trace(args, n_batches):
    from b200_router import traces
    return traces.SharedPrefixTrace(n_total=n_batches * args.batch, batch=args.batch, n_sys=N_SYS,
                                    sys_chars=SYS_CHARS, suffix_lo=SUFFIX[0], suffix_hi=SUFFIX[1],
                                    seed=SEED)


# ---------------------------------------------------------------- clocks
class ClockSampler:
    """SM clock and throttle reasons during the timed region, read through NVML.

    NVML queries take a driver lock and stall kernel launches of every process on the box
    (measured: a background poller more than doubled the multi-GPU step time), so nothing
    polls while the host is still enqueueing.  The timed steps are enqueued first; then,
    while the GPU is still executing them, the main thread samples until the closing event
    has completed: the samples fall inside the timed region without perturbing it."""

    def __init__(self, gpu_index, enabled=True):
        self.gpu = gpu_index
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self.h = None
        if not enabled:                # only one process per box talks to NVML
            return
        try:
            import pynvml as nv
            self.nv = nv
            nv.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            idx = gpu_index
            if vis and all(x.strip().isdigit() for x in vis.split(",")):
                idx = int(vis.split(",")[gpu_index])
            self.h = nv.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
        except Exception:
            self.h = None

    def sample(self):
        if self.h is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        try:
            self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            for bit, nm in names.items():
                if r & bit:
                    self.reasons.add(nm)
        except Exception:
            self.h = None

    def sample_until(self, event, limit=4):
        """A few samples while `event` (recorded after the timed steps) has not completed."""
        self.sample()
        n0 = len(self.samples)
        while not event.query() and len(self.samples) - n0 < limit:
            time.sleep(0.0003)
            if not event.query():
                self.sample()

    def result(self, reset=False):
        out = {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
               "samples": len(self.samples)}
        if self.samples:
            out["sm_mhz"] = float(np.median(self.samples))
        if reset:
            self.samples, self.reasons = [], set()
        return out


_SAMPLER = None


def sampler(local=0, enabled=True):
    """One NVML handle per process (start-up takes milliseconds: never inside a timed region)."""
    global _SAMPLER
    if _SAMPLER is None:
        _SAMPLER = ClockSampler(local, enabled)
        _SAMPLER.sample()
        _SAMPLER.samples.clear()
    return _SAMPLER


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return json.load(f), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md)"


def load_ncu_traffic(kernel):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture, with the
    commit the capture was taken at (profiles/ncu_traffic.json); None when there is none."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            d = json.load(f)
        return d[kernel]["dram_bytes_per_launch"], d.get("commit")
    except Exception:
        return None, None


# ------------------------------------------------------- CPU (oracle) legs
def prompts_of(buf, starts, lens, count):
    out = []
    for i in range(min(count, len(starts))):
        s = int(starts[i])
        out.append(buf[s:s + int(lens[i])].tobytes().decode("ascii"))
    return out


def cpu_port_rate(prompts, avail, n_endpoints=N_ENDPOINTS, seed=SEED):
    """Python restatement of the reference router, one asyncio loop, 1 core.
    Returns (decisions/s, decisions as endpoint indices)."""
    from oracle import prefix_ref
    from oracle.chooser import ChoiceShim
    from b200_router import traces
    urls = traces.endpoint_urls(n_endpoints)
    index_of = {u: i for i, u in enumerate(urls)}
    router = prefix_ref.PrefixRouterRef(ChoiceShim(index_of, seed))
    eps = [urls[i] for i in range(n_endpoints) if (avail >> i) & 1]
    t0 = time.perf_counter()
    got = router.route_many(eps, prompts)
    dt = time.perf_counter() - t0
    return len(prompts) / dt, [index_of[u] for u in got]


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import asyncio
    from oracle import prefix_ref
    from oracle.chooser import ChoiceShim
    from b200_router import traces
    sample = 8192                      # requests per step: bounded sample of a batch
    nb = args.warmup + args.steps
    tr = make_trace(args, nb)
    urls = traces.endpoint_urls(N_ENDPOINTS)
    index_of = {u: i for i, u in enumerate(urls)}
    router = prefix_ref.PrefixRouterRef(ChoiceShim(index_of, SEED))
    steps = []
    for b in range(nb):
        buf, starts, lens = tr.batch_arrays(b)
        steps.append((avail_mask(b), prompts_of(buf, starts, lens, sample)))

    async def one(avail, prompts):
        aset = set(urls[i] for i in range(N_ENDPOINTS) if (avail >> i) & 1)
        for p in prompts:
            await router.route_prompt(aset, p)

    for b in range(args.warmup):
        asyncio.run(one(*steps[b]))
    t0 = time.perf_counter()
    nreq = nbytes = 0
    for b in range(args.warmup, nb):
        asyncio.run(one(*steps[b]))
        nreq += len(steps[b][1])
        nbytes += sum(len(p) for p in steps[b][1])
    dt = time.perf_counter() - t0
    value = nreq / dt
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": workload_config(args, args.gpus),
            "prompt_gbs": nbytes / dt / 1e9,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": "port",
                             "sample": "first %d requests of each of %d batches; Python "
                                       "restatement of the reference (oracle/prefix_ref.py), one "
                                       "asyncio loop; host has %d cpus" % (sample, args.steps,
                                                                           os.cpu_count())},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ------------------------------------------------------------- B200 arm
class Ctx:
    """Process-wide set-up shared by the workloads of one invocation."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device (the B200 path has no CPU fallback)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1 and not dist.is_initialized():
            dist.init_process_group("nccl", device_id=self.dev)
        # NVML is initialised here, well before any timed region: its start-up takes
        # milliseconds on the host and every rank waits for the slowest one
        self.sampler = sampler(self.local, enabled=(self.rank == 0))

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x, dtype=None):
        if self.world == 1:
            return x
        t = self.torch.tensor([x], device=self.dev, dtype=dtype or self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.item()

    def sum_over_ranks(self, x):
        if self.world == 1:
            return x
        t = self.torch.tensor([x], device=self.dev, dtype=self.torch.int64)
        self.dist.all_reduce(t)
        return int(t.item())


def new_router(cx, cap_log2, max_requests, max_chunks):
    if cx.world == 1:
        # public single-GPU API: table + two-stream batch pipeline (hash stage of batch
        # b+1 overlaps the route stage of batch b; results stay sequential)
        from b200_router.trie import GpuHashTrie, RoutePipeline
        pipe = RoutePipeline(GpuHashTrie(cap_log2=cap_log2, device=cx.dev),
                             depth=int(os.environ.get("BENCH_PIPE_DEPTH", "3")))
        pipe.reserve(max_requests, max_chunks)
        return pipe
    # multi-GPU: table partitioned by first-chunk hash, chunk hashes and results shipped
    # over NVLink peer memory by the library's own kernels (the variant that scales).
    # BENCH_EXCHANGE=replicated selects the replicated-table variant (NCCL all-gather of
    # the chunk hashes, every rank routes the global batch) for comparison.
    from b200_router.distributed import PartitionedPrefixRouter, ShardedPrefixRouter
    cls = {"nccl": ShardedPrefixRouter, "replicated": ShardedPrefixRouter}.get(
        os.environ.get("BENCH_EXCHANGE", ""), PartitionedPrefixRouter)
    return cls(cap_log2=cap_log2, device=cx.dev, max_requests=max_requests, max_chunks=max_chunks)


def clear_router(cx, router):
    """Fresh table (and, multi-GPU, every rank past its last batch) between passes."""
    cx.barrier()
    (router.trie if hasattr(router, "trie") else router.ops.trie).clear()
    cx.barrier()


def route_any(cx, router, p, avail, seq, out=None, ready=True):
    if cx.world == 1:
        return router.route(p, avail, SEED, seq, out=out, inputs_ready=ready)
    if hasattr(router, "pipelined"):                       # partitioned: pipelined
        return router.route(p, avail, SEED, seq, out=out, inputs_ready=ready)
    return router.route(p, avail, SEED, seq, out=out)


def count_nodes(cx, router):
    st = router.check_status(count_nodes=True)
    return cx.sum_over_ranks(int(st["num_nodes"])), st


def parity_check(cx, router, seq, per_rank=2048, n_batches=4):
    """Churn-bearing batches on a cleared table through the SAME router object the timed
    passes used: per rank, decisions / depths / candidate sets against the C oracle routing
    the rank-major concatenation sequentially; at the end the union of the ranks' tables
    against the oracle's trie.  Returns (ok, detail, seq)."""
    from b200_router import traces
    from b200_router.packing import pack_arrays
    from oracle import coracle
    torch, dist = cx.torch, cx.dist
    W = cx.world
    tr = traces.SharedPrefixTrace(n_total=n_batches * W * per_rank, batch=per_rank, n_sys=16,
                                  sys_chars=1024, suffix_lo=1, suffix_hi=700, seed=SEED + 1)
    clear_router(cx, router)
    orc = coracle.TrieOracle() if cx.rank == 0 else None
    ok, detail = True, []
    for g in range(n_batches):
        avail = 0xFF & ~(1 << 2) if g in (1, 2) else 0xFF
        buf, starts, lens = tr.batch_arrays(g * W + cx.rank)
        ep, depth, matched = route_any(cx, router, pack_arrays(buf, starts, lens, pin=False).to(cx.dev),
                                       avail, seq, ready=False)
        got = (ep.cpu().numpy().copy(), depth.cpu().numpy().copy(),
               matched.cpu().numpy().view(np.uint64).copy())
        if W > 1:
            allgot = [None] * W
            dist.all_gather_object(allgot, got)
        else:
            allgot = [got]
        if cx.rank == 0:
            off = 0
            for r in range(W):
                b2, s2, l2 = tr.batch_arrays(g * W + r)
                want = orc.route_batch(b2, s2, l2, avail, SEED, seq + off)
                off += len(s2)
                for k, nm in enumerate(("endpoint", "depth", "candidates")):
                    if not (allgot[r][k] == want[k]).all():
                        ok = False
                        detail.append("batch %d rank %d: %s differ" % (g, r, nm))
        seq += W * per_rank
    mine = router.dump() if hasattr(router, "dump") else router.trie.dump()
    if W > 1:
        dumps = [None] * W
        dist.all_gather_object(dumps, mine)
    else:
        dumps = [mine]
    if cx.rank == 0:
        replicated = W > 1 and not hasattr(router, "pipelined")
        merged = {}
        for d in dumps:
            for path, m in d.items():
                if path == ():
                    merged[path] = merged.get(path, 0) | m
                elif replicated:
                    if merged.setdefault(path, m) != m:
                        ok = False
                        detail.append("replicas differ")
                else:
                    if path in merged:
                        ok = False
                        detail.append("a node lives in two partitions")
                    merged[path] = m
        if merged != orc.dump():
            ok = False
            detail.append("final trie differs from the oracle's")
    if W > 1:
        t = torch.tensor([1 if ok else 0], device=cx.dev)
        dist.broadcast(t, 0)
        ok = bool(t.item())
    return ok, detail, seq


def run_b200(args):
    cx = Ctx()
    torch = cx.torch
    from b200_router import _lib
    from b200_router.packing import pack_arrays
    world, rank, dev = cx.world, cx.rank, cx.dev
    W, K, R = args.warmup, args.steps, max(args.passes, 1)
    nb = W + K
    if os.environ.get("BENCH_HASH_MODE"):
        _lib.check(_lib.lib().b200r_set_hash_mode(int(os.environ["BENCH_HASH_MODE"])))
    if os.environ.get("BENCH_CLAIM_MODE"):      # tuning experiments only (see include/b200_router.h)
        _lib.check(_lib.lib().b200r_set_claim_mode(int(os.environ["BENCH_CLAIM_MODE"])))

    # ---- synthetic batches (host, pinned), this rank's shard of every global batch
    tr = make_trace(args, nb * world)
    host = []
    for b in range(nb):
        buf, starts, lens = tr.batch_arrays(b * world + rank)
        host.append(pack_arrays(buf, starts, lens, pin=True))
    max_chunks = max(p.max_chunks for p in host)
    if not args.no_extras:
        max_chunks = max(max_chunks, 8192 * 256 + 1024)      # the C4 extra shares the router
    nodes_est = nb * world * args.batch * 8
    cap_log2 = max(16, min(28, int(np.ceil(np.log2(nodes_est * 1.3)))))
    out = (torch.empty(args.batch, dtype=torch.int32, device=dev),
           torch.empty(args.batch, dtype=torch.int32, device=dev),
           torch.empty(args.batch, dtype=torch.int64, device=dev))
    resident = [p.to(dev, non_blocking=False) for p in host]
    router = new_router(cx, cap_log2, args.batch, max_chunks)
    seq = 0                                    # global tie-break counter (never reset)

    # ---- pass 1: device-resident batches; R passes of (W warm-up + K timed steps), each on
    #      a cleared table; every timed region bracketed by events, barrier and synchronize
    per_pass, launches, host_submit_ms = [], 0, 0.0
    nodes_before = nodes_after = 0
    for r in range(R):
        clear_router(cx, router)
        for b in range(W):
            route_any(cx, router, resident[b], avail_mask(b), seq, out if world > 1 else None)
            seq += args.batch * world
        cx.barrier()
        if r == 0:
            nodes_before, _ = count_nodes(cx, router)
            cx.barrier()
        launches0 = _lib.launch_count()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        th0 = time.perf_counter()
        for b in range(W, nb):
            route_any(cx, router, resident[b], avail_mask(b), seq, out if world > 1 else None)
            seq += args.batch * world
        host_submit_ms += (time.perf_counter() - th0) * 1e3
        ev1.record()
        if rank == 0:
            cx.sampler.sample_until(ev1)     # GPU is still inside the timed region here
        cx.barrier()
        launches += _lib.launch_count() - launches0
        per_pass.append(cx.max_over_ranks(ev0.elapsed_time(ev1)) / K)
        if r == 0:
            nodes_after, st = count_nodes(cx, router)
    clocks = cx.sampler.result(reset=True)
    ms_step = float(np.mean(per_pass))
    n_dec_step = args.batch * world
    payload_step = cx.sum_over_ranks(sum(p.payload_bytes for p in host[W:])) / K
    chunks_step = cx.sum_over_ranks(sum(p.max_chunks for p in host[W:])) / K
    new_step = (nodes_after - nodes_before) / K
    value = n_dec_step / (ms_step * 1e-3)
    whole_bytes = payload_step + 8 * n_dec_step + 32 * chunks_step + 32 * new_step + 4 * n_dec_step

    # ---- pass 2: the same W+K batches on a cleared table, with per-kernel CUDA events
    # around every launch of the K timed steps (roofline of the longest kernel, kernel shares).
    # Single stream here (no batch overlap), so that every kernel is timed on its own and
    # its share of the step is comparable with the serialised ncu launch list.
    clear_router(cx, router)
    serial = router.trie if world == 1 else router
    for b in range(W):
        if world == 1:
            serial.route(resident[b], avail_mask(b), SEED, seq, out=out)
        else:
            route_any(cx, router, resident[b], avail_mask(b), seq, out)
        seq += args.batch * world
    cx.barrier()
    _lib.prof_enable(True)
    for b in range(W, nb):
        if world == 1:
            serial.route(resident[b], avail_mask(b), SEED, seq, out=out)
        else:
            route_any(cx, router, resident[b], avail_mask(b), seq, out)
        seq += args.batch * world
    torch.cuda.synchronize()
    prof = _lib.prof_read()
    _lib.prof_enable(False)
    cx.barrier()
    my_payload = sum(p.payload_bytes for p in host[W:])
    my_tail = sum(int((p.lens.numpy() % 128).sum()) for p in host[W:])
    my_chunks = sum(p.max_chunks for p in host[W:])
    my_new = (nodes_after - nodes_before) / world
    # algorithmic bytes of each kernel class over the K timed launches of this rank (SURVEY 8d:
    # prompt bytes once, 8 B offset, one 32-B probe per depth, 32 B per new node, 4 B decision;
    # the ragged last chunks are hashed by the prep kernel and credited there)
    algo = {"hash": my_payload - my_tail,
            "prep": my_tail + 8 * K * args.batch,
            "claim": 32 * my_chunks + 32 * my_new,
            "select": 4 * K * args.batch,
            "commit": 4 * K * args.batch}
    peaks, peak_src = load_peaks()
    kern_ms = {k: v[0] / K for k, v in prof.items() if v[1]}
    prof_total = sum(v[0] for v in prof.values())
    core = {k: prof[k] for k in ("prep", "hash", "claim", "select", "commit") if prof[k][1]}
    if "commit" not in core:                  # select + commit ran as one kernel (k_select_commit)
        algo["select"] += algo["commit"]
    dom = max(core, key=lambda k: core[k][0])
    dom_ms, dom_n = core[dom]
    from b200_router.trie import claim_kernel_name
    kernel_name = {"prep": "k_prep_scan", "hash": "k_hash_chunks",
                   "claim": claim_kernel_name(args.batch, my_chunks // max(K, 1)),
                   "select": "k_select_commit" if "commit" not in core else "k_select",
                   "commit": "k_commit"}[dom]
    achieved = algo[dom] / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0
    traffic, traffic_commit = load_ncu_traffic(kernel_name)
    per_kernel = {k: {"ms_per_step": core[k][0] / K,
                      "algorithmic_gbs": algo[k] / (core[k][0] * 1e-3) / 1e9,
                      "frac": algo[k] / (core[k][0] * 1e-3) / 1e9 / peaks["hbm_gbs"]}
                  for k in core if core[k][0] > 0}
    roofline = {"bound": "hbm", "kernel": kernel_name, "achieved": achieved, "peak": peaks["hbm_gbs"],
                "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"], "traffic": traffic,
                "traffic_source": None if traffic is None else
                "profiles/ncu_traffic.json, ncu --set full at commit %s" % traffic_commit,
                "peak_source": peak_src, "launches": dom_n, "ms_per_launch": dom_ms / max(dom_n, 1),
                "share_of_step": dom_ms / prof_total if prof_total else None,
                "kernel_ms_per_step": kern_ms, "per_kernel": per_kernel,
                "whole_path": {"bytes_per_step": whole_bytes,
                               "formula": "sum(L + 8 + 32*C + 32*C_new + 4) per request (SURVEY 8d)",
                               "chunks_per_step": chunks_step, "new_nodes_per_step": new_step,
                               "achieved": whole_bytes / (ms_step * 1e-3) / 1e9,
                               "frac": whole_bytes / (ms_step * 1e-3) / 1e9 / peaks["hbm_gbs"],
                               "prompt_only_frac": payload_step / (ms_step * 1e-3) / 1e9 / peaks["hbm_gbs"]}}

    # ---- parity gate: same router object, cleared table
    parity, parity_detail, seq = parity_check(cx, router, seq)

    # ---- pass 3: end to end from pinned host buffers through the public API.  Every step
    # copies its packed batch host->device and its decisions device->host inside the timed
    # region; the host reads step b's decisions after submitting step b+1 (the pipeline
    # overlaps the copy of the next batch with the routing of the current one).
    resident = None
    torch.cuda.empty_cache()
    clear_router(cx, router)
    res_host = [torch.empty(args.batch, dtype=torch.int32).pin_memory() for _ in range(3)]
    res_ev = [torch.cuda.Event() for _ in range(3)]

    def submit(b, seq, j):
        if world == 1:
            ep, _, _ = router.route(host[b], avail_mask(b), SEED, seq)
        elif hasattr(router, "pipelined"):      # host inputs are copied on the send stream
            ep, _, _ = router.route(host[b], avail_mask(b), SEED, seq, out=out)
        else:
            ep, _, _ = router.route(host[b].to(dev, non_blocking=True), avail_mask(b), SEED, seq,
                                    out=out)
        res_host[j][:ep.numel()].copy_(ep, non_blocking=True)
        res_ev[j].record()

    checksum = 0
    for b in range(W):
        submit(b, seq, b % 3)
        seq += args.batch * world
    cx.barrier()
    t0 = time.perf_counter()
    prev = None
    for b in range(W, nb):
        j = b % 3
        submit(b, seq, j)
        seq += args.batch * world
        if prev is not None:
            res_ev[prev].synchronize()
            checksum += int(res_host[prev][0]) + int(res_host[prev][-1])   # host reads the result
        prev = j
    res_ev[prev].synchronize()
    checksum += int(res_host[prev][0]) + int(res_host[prev][-1])
    cx.barrier()
    e2e_s = cx.max_over_ranks(time.perf_counter() - t0)
    h2d = int(np.mean([p.h2d_bytes() for p in host[W:]]))
    e2e = {"value": K * n_dec_step / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": args.batch * 4, "ms_per_step": e2e_s / K * 1e3,
           "h2d_gbs_per_gpu": h2d * K / e2e_s / 1e9}

    # ---- the same from Python str objects (what a caller of route_batch holds): native packer
    #      (libb200pack.so) into a pinned arena + H2D + route + D2H, pipelined.  Host-bound:
    #      reported next to e2e, not instead of it.
    e2e_strings = None
    if rank == 0 and world == 1 and not args.no_cpu:
        from b200_router.packing import PackArena, PackPipeline
        ns = min(8, nb)
        strs = [prompts_of(*tr.batch_arrays(b), args.batch) for b in range(ns)]
        workers = int(os.environ.get("BENCH_PACK_WORKERS", "3"))
        ahead = workers
        pp = PackPipeline(arena=PackArena(slots=ahead + 3), workers=workers)   # packs `ahead` batches ahead
        for rep in range(3):                      # reps 0-1 touch every arena slot (pinned allocation)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            prev = None
            for k in range(min(ahead, ns)):
                pp.submit(strs[k])
            for k in range(ns):
                pk = pp.next()
                if k + ahead < ns:
                    pp.submit(strs[k + ahead])
                ep, _, _ = router.route(pk, avail_mask(W), SEED, seq)
                seq += args.batch
                res_host[k % 3][:ep.numel()].copy_(ep, non_blocking=True)
                res_ev[k % 3].record()
                if prev is not None:
                    res_ev[prev].synchronize()
                prev = k % 3
            res_ev[prev].synchronize()
            str_s = time.perf_counter() - t0
        pp.close()
        e2e_strings = {"value": ns * args.batch / str_s, "unit": UNIT, "ms_per_step": str_s / ns * 1e3,
                       "note": "input = list of Python str per step; native packer on %d worker threads "
                               "(copy phase without the GIL, %d cpus) + H2D + route + D2H" %
                               (workers, os.cpu_count())}
        del strs

    # ---- CPU baseline (rank 0, N=1 only): the Python port on the first requests of timed
    #      batch 0, and the same sample through the GPU path on an empty table
    cpu = None
    port_ok = None
    if rank == 0 and world == 1 and not args.no_cpu:
        from b200_router.trie import GpuHashTrie
        b = W
        cnt = min(args.cpu_sample, args.batch)
        buf, starts, lens = tr.batch_arrays(b)
        prompts = prompts_of(buf, starts, lens, cnt)
        rate, want = cpu_port_rate(prompts, avail_mask(b))
        t2 = GpuHashTrie(cap_log2=max(16, int(np.ceil(np.log2(cnt * 24 * 1.2)))), device=dev)
        ep, _, _ = t2.route(pack_arrays(buf, starts[:cnt], lens[:cnt], pin=False),
                            avail_mask(b), SEED, 0)
        port_ok = bool(ep.cpu().tolist() == want)
        t2.close()
        cpu = {"value": rate, "unit": UNIT, "cores": 1, "kind": "port",
               "sample": "first %d requests of timed batch 0 on an empty table; Python "
                         "restatement of the reference (oracle/prefix_ref.py), one asyncio loop, "
                         "host has %d cpus" % (cnt, os.cpu_count()),
               "gpu_decisions_equal": port_ok}

    # ---- other configs of BASELINE.json, shorter, in the same record
    extras = None
    if not args.no_extras:
        extras = run_extras(cx, args, router, seq)

    if rank == 0:
        line = {"metric": METRIC, "value": value if parity and port_ok is not False else None,
                "unit": UNIT, "n_gpus": world, "steps": K,
                "warmup": W, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u64", "data": "synthetic",
                "config": workload_config(args, world),
                "ms_per_step_by_pass": per_pass, "timed_ms_total": ms_step * K * R,
                "prompt_gbs": payload_step / (ms_step * 1e-3) / 1e9,
                "prompt_frac_of_hbm": payload_step / (ms_step * 1e-3) / 1e9 / peaks["hbm_gbs"],
                "clocks": clocks, "host_submit_ms_per_step": host_submit_ms / (K * R),
                "e2e": e2e, "e2e_strings": e2e_strings,
                "gpu_launches": int(launches // R),
                "roofline": roofline,
                "cpu_baseline": cpu, "parity_sample_ok": parity, "parity_detail": parity_detail,
                "parity_sample": "4 global batches of %d requests (2048 per rank), endpoint 2 down in "
                                 "batches 1-2, cleared table: decisions, depths, candidate sets per rank "
                                 "and the final trie (union of the ranks' tables) vs the C oracle" %
                                 (2048 * world),
                "table": {"nodes": st["num_nodes"], "capacity": st["capacity"]},
                "extras": extras}
        print(json.dumps(line))
    if world > 1:
        cx.dist.destroy_process_group()
    if not parity:
        sys.exit(1)


# ----------------------------------------------------------- other configs
def timed_region(cx, fn, steps):
    """`steps` calls of fn(i) bracketed by barrier + synchronize and CUDA events; clocks sampled
    by rank 0 while the GPU is still busy.  Returns (ms, clocks), ms = max over ranks."""
    torch = cx.torch
    cx.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(steps):
        fn(i)
    ev1.record()
    if cx.rank == 0:
        cx.sampler.sample_until(ev1)
    cx.barrier()
    return cx.max_over_ranks(ev0.elapsed_time(ev1)), cx.sampler.result(reset=True)


def session_line(cx, args, K=None, W=None):
    """configs[1]: 1M synthetic session ids over 16 endpoints, batched ring lookup on 1 GPU.
    ALU-bound (one MD5 block per key); the HBM fraction is informational."""
    torch = cx.torch
    from b200_router import traces
    from b200_router.packing import pack_arrays
    from b200_router.ring import GpuHashRing
    n = 1 << 20
    K = K or args.steps
    W = W or args.warmup
    nodes = traces.endpoint_urls(16)
    ring = GpuHashRing(nodes, device=cx.dev)
    buf, starts, lens = traces.session_ids_packed(n, seed=0)
    host = pack_arrays(buf, starts, lens, pin=True)
    res = host.to(cx.dev, non_blocking=False)
    for _ in range(W):
        ring.lookup_packed(res)
    ms, clocks = timed_region(cx, lambda i: ring.lookup_packed(res), K)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(K):
        ring.lookup_packed(host.to(cx.dev, non_blocking=True)).cpu()
    e2e_s = time.perf_counter() - t0
    # membership change (routing_logic.py:84-103): remove + re-add one endpoint, next lookup
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ring.remove_node(nodes[3])
    ring.add_node(nodes[3])
    ring.lookup_packed(res)
    torch.cuda.synchronize()
    rebuild_ms = (time.perf_counter() - t0) * 1e3 - ms / K
    cpu = None
    if not args.no_cpu:
        from oracle import ring_ref
        ref = ring_ref.RingRef(nodes)
        ids = traces.session_ids_uuid(100000, 0)
        t0 = time.perf_counter()
        want = [ref.get_node(k) for k in ids]
        dt = time.perf_counter() - t0
        cpu = {"value": len(ids) / dt, "unit": "lookups/s", "cores": 1, "kind": "port",
               "sample": "100000 uuid ids through oracle/ring_ref.py (restated uhashring ring; "
                         "parity unpinned: the package is not installable here)",
               "gpu_decisions_equal": ring.get_node_batch(ids[:5000]) == want[:5000]}
    ring.close()
    peaks, _ = load_peaks()
    algo = n * (36 + 8 + 4 + 4)
    return {"metric": "session_route_lookups_per_sec", "value": n * K / (ms * 1e-3), "unit": "lookups/s",
            "n_gpus": 1, "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": "C2: 1M uuid-shaped session ids, 16 endpoints, ring of 2560 points"},
            "clocks": clocks,
            "e2e": {"value": n * K / e2e_s, "unit": "lookups/s", "h2d_bytes_per_step": host.h2d_bytes(),
                    "d2h_bytes_per_step": n * 4},
            "roofline": {"bound": "hbm", "kernel": "k_ring_lookup", "achieved": algo / (ms / K * 1e-3) / 1e9,
                         "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": algo / (ms / K * 1e-3) / 1e9 / peaks["hbm_gbs"], "traffic": None,
                         "note": "ALU-bound by construction: one 64-step MD5 block per 36-byte key"},
            "md5_blocks_per_sec": n * K / (ms * 1e-3),
            "ring_rebuild_ms": rebuild_ms, "cpu_baseline": cpu, "gpu_launches": K}


def mrqa_line(cx, args, router, seq, users=5000, rounds=10, label="C5"):
    """configs[4]: replay of a benchmarks/multi-round-qa style trace (users x rounds, every
    round carries the whole chat history) through the GPU router, one global batch per round
    (users split over the ranks), vs the Python restatement of the reference on a sample."""
    torch = cx.torch
    from b200_router import traces
    from b200_router.packing import pack_prompts
    W = cx.world
    per = users // W
    prompts, _ = traces.multi_round_qa(users=per * W, rounds=rounds, sys_len=1000, hist_len=200,
                                       answer_len=100)
    nu = per * W
    mine = [prompts[r * nu + cx.rank * per: r * nu + (cx.rank + 1) * per] for r in range(rounds)]
    host = [pack_prompts(m) for m in mine]
    resident = [p.to(cx.dev, non_blocking=False) for p in host]
    payload = cx.sum_over_ranks(sum(p.payload_bytes for p in host))
    avail = (1 << 4) - 1
    out = (torch.empty(per, dtype=torch.int32, device=cx.dev),
           torch.empty(per, dtype=torch.int32, device=cx.dev),
           torch.empty(per, dtype=torch.int64, device=cx.dev))
    state = {"seq": seq}

    def step(i):
        route_any(cx, router, resident[i], avail, state["seq"], out if W > 1 else None)
        state["seq"] += nu

    for _ in range(max(args.warmup, 1)):
        clear_router(cx, router)
        for i in range(rounds):
            step(i)
    clear_router(cx, router)
    ms, clocks = timed_region(cx, step, rounds)
    nodes, _ = count_nodes(cx, router)
    clear_router(cx, router)
    cx.barrier()
    t0 = time.perf_counter()
    last = None
    for i in range(rounds):
        last = route_any(cx, router, host[i], avail, state["seq"], out if W > 1 else None, ready=False)
        state["seq"] += nu
    last[0].cpu()
    cx.barrier()
    e2e_s = cx.max_over_ranks(time.perf_counter() - t0)
    cpu = None
    if cx.rank == 0 and W == 1 and not args.no_cpu:
        from b200_router.trie import GpuHashTrie
        sample = prompts[:nu * 2]                    # the first two rounds
        rate, want = cpu_port_rate(sample, avail, n_endpoints=4)
        t = GpuHashTrie(cap_log2=20, device=cx.dev)
        ep0, _, _ = t.route(host[0].to(cx.dev), avail, SEED, 0)
        ep1, _, _ = t.route(host[1].to(cx.dev), avail, SEED, nu)
        ok = (ep0.cpu().tolist() + ep1.cpu().tolist()) == want
        t.close()
        cpu = {"value": rate, "unit": UNIT, "cores": 1, "kind": "port", "gpu_decisions_equal": ok,
               "sample": "first 2 rounds (%d requests) through oracle/prefix_ref.py" % len(sample)}
    peaks, _ = load_peaks()
    n = nu * rounds
    return {"metric": METRIC, "value": n / (ms * 1e-3), "unit": UNIT, "n_gpus": W, "steps": rounds,
            "warmup": max(args.warmup, 1), "ms_per_step": ms / rounds, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": "%s: multi-round-QA replay, %d users x %d rounds, 4 endpoints, one "
                                   "global batch per round, prompts %.1f KB on average" %
                                   (label, nu, rounds, payload / n / 1e3)},
            "clocks": clocks, "prompt_gbs": payload / (ms * 1e-3) / 1e9,
            "prompt_frac_of_hbm": payload / (ms * 1e-3) / 1e9 / peaks["hbm_gbs"],
            "e2e": {"value": n / e2e_s, "unit": UNIT,
                    "h2d_bytes_per_step": sum(p.h2d_bytes() for p in host) // rounds,
                    "d2h_bytes_per_step": per * 4},
            "cpu_baseline": cpu, "table": {"nodes": nodes}}, state["seq"]


def c4_line(cx, args, router, seq, fracs=(0.0, 0.5, 0.94), steps=3, per_rank=8192):
    """configs[3]: 32768-char prompts (256 chunks), 64 endpoints, groups of 64 prompts sharing
    the leading `frac` of their chunks, per_rank requests per GPU and step, global order =
    rank-major.  Reports per shared fraction the step time and, multi-GPU, what crosses NVLink."""
    torch = cx.torch
    from b200_router import _lib, traces
    from b200_router.packing import pack_arrays
    W = cx.world
    full = (1 << 64) - 1
    out = (torch.empty(per_rank, dtype=torch.int32, device=cx.dev),
           torch.empty(per_rank, dtype=torch.int32, device=cx.dev),
           torch.empty(per_rank, dtype=torch.int64, device=cx.dev))
    state = {"seq": seq}
    res = {}
    peaks, _ = load_peaks()
    for frac in fracs:
        nb = 1 + steps
        resident = []
        for b in range(nb):
            buf, starts, lens = traces.long_context_batch(n=per_rank, chars=32768, group=64,
                                                          shared_frac=frac,
                                                          seed=1000 * (b * W + cx.rank) + int(frac * 100))
            resident.append(pack_arrays(buf, starts, lens, pin=False).to(cx.dev, non_blocking=False))
        clear_router(cx, router)

        def step(i):
            route_any(cx, router, resident[i], full, state["seq"], out if W > 1 else None)
            state["seq"] += per_rank * W

        step(0)
        n0, _ = count_nodes(cx, router)
        ms, clocks = timed_region(cx, lambda i: step(1 + i), steps)
        n1, _ = count_nodes(cx, router)
        # per-kernel-class times of one more pass (exchange = push + wait classes)
        clear_router(cx, router)
        step(0)
        cx.barrier()
        _lib.prof_enable(True)
        for i in range(steps):
            step(1 + i)
        torch.cuda.synchronize()
        prof = _lib.prof_read()
        _lib.prof_enable(False)
        cx.barrier()
        chunks = per_rank * 256
        payload = per_rank * 32768 * W
        line = {"ms_per_step": ms / steps, "decisions_per_s": per_rank * W * steps / (ms * 1e-3),
                "prompt_gbs": payload * steps / (ms * 1e-3) / 1e9,
                "prompt_frac_of_hbm_per_gpu": payload / W * steps / (ms * 1e-3) / 1e9 / peaks["hbm_gbs"],
                "new_nodes_per_step": (n1 - n0) / steps, "clocks": clocks,
                "kernel_ms_per_step": {k: v[0] / steps for k, v in prof.items() if v[1]}}
        if W > 1:
            # every request's chunk hashes (8 B per chunk) + {index, count} word go to the rank
            # owning its first chunk, {endpoint, depth, candidates} (16 B) come back
            line["exchange"] = {
                "bytes_per_rank_per_step": int((chunks * 8 + per_rank * 24) * (W - 1) / W),
                "note": "expected share leaving the rank, (W-1)/W of hashes + request words + results",
                "ms_per_step_push_plus_wait": (prof["push"][0] + prof["wait"][0]) / steps}
        res["shared_%.2f" % frac] = line
        resident = None
        torch.cuda.empty_cache()
    return {"config": {"workload": "C4: 32768-char prompts (256 chunks), 64 endpoints, %d requests per GPU "
                                   "and step, groups of 64 sharing the leading fraction of their chunks"
                                   % per_rank, "n_gpus": W, "steps": steps, "warmup": 1},
            "by_shared_fraction": res}, state["seq"]


def latency_line(cx, args):
    """The reference calls route_request once per HTTP request (request.py:252-257).  p50 / p99
    of one `await router.route_request(...)` at a time, and decisions/s at several levels of
    concurrency through the micro-batching front, next to the Python restatement of the
    reference on the same prompts (1 core)."""
    import asyncio
    from b200_router import routers, traces
    tr = traces.SharedPrefixTrace(n_total=8192, batch=8192, n_sys=64, sys_chars=SYS_CHARS,
                                  suffix_lo=SUFFIX[0], suffix_hi=SUFFIX[1], seed=SEED + 7)
    prompts = prompts_of(*tr.batch_arrays(0), 8192)
    urls = traces.endpoint_urls(N_ENDPOINTS)

    class Ep:
        def __init__(self, url):
            self.url = url
    eps = [Ep(u) for u in urls]
    out = {}
    routers._reset_singletons_for_tests()
    r = routers.PrefixAwareRouter(cap_log2=20, seed=SEED, device=cx.dev)

    async def serial(ps):
        lat = []
        for p in ps:
            t0 = time.perf_counter()
            await r.route_request(eps, None, None, None, {"prompt": p})
            lat.append(time.perf_counter() - t0)
        return lat
    asyncio.run(serial(prompts[:64]))                       # warm-up
    lat = np.array(asyncio.run(serial(prompts[64:1088]))) * 1e6
    out["one_at_a_time"] = {"p50_us": float(np.percentile(lat, 50)), "p99_us": float(np.percentile(lat, 99)),
                            "decisions_per_s": 1e6 / float(lat.mean()), "requests": len(lat)}
    # the same through the batch kernels (a batch of one: six launches, two copies) -- what the
    # latency path (b200r_prefix_route_one: one launch, prompt read from pinned host memory)
    # replaces
    lat1 = []
    for p in prompts[1088:1600]:
        t0 = time.perf_counter()
        r.route_batch(eps, [p])
        lat1.append(time.perf_counter() - t0)
    lat1 = np.array(lat1[64:]) * 1e6
    out["one_at_a_time_as_batch_of_one"] = {"p50_us": float(np.percentile(lat1, 50)),
                                            "p99_us": float(np.percentile(lat1, 99))}
    routers._reset_singletons_for_tests()
    conc = {}
    for c in (8, 64, 512, 4096):
        routers._reset_singletons_for_tests()
        rb = routers.PrefixAwareRouter(cap_log2=20, seed=SEED, device=cx.dev, micro_batch_ms=0.2,
                                       micro_batch_max=c)

        async def burst(ps):
            return await asyncio.gather(*[rb.route_request(eps, None, None, None, {"prompt": p})
                                          for p in ps])

        async def run():
            await burst(prompts[:c])                         # warm-up
            t0 = time.perf_counter()
            n = 0
            for k in range(0, 4096, c):
                await burst(prompts[k:k + c])
                n += c
            return n / (time.perf_counter() - t0)
        conc[str(c)] = asyncio.run(run())
    routers._reset_singletons_for_tests()
    out["micro_batched_decisions_per_s_by_concurrency"] = conc
    if not args.no_cpu:
        from oracle import prefix_ref
        from oracle.chooser import ChoiceShim
        index_of = {u: i for i, u in enumerate(urls)}
        ref = prefix_ref.PrefixRouterRef(ChoiceShim(index_of, SEED))

        async def ref_serial(ps):
            lat = []
            aset = set(urls)
            for p in ps:
                t0 = time.perf_counter()
                await ref.route_prompt(aset, p)
                lat.append(time.perf_counter() - t0)
            return lat
        asyncio.run(ref_serial(prompts[:64]))
        rl = np.array(asyncio.run(ref_serial(prompts[64:1088]))) * 1e6
        out["reference_port_one_at_a_time"] = {"p50_us": float(np.percentile(rl, 50)),
                                               "p99_us": float(np.percentile(rl, 99)),
                                               "decisions_per_s": 1e6 / float(rl.mean())}
        per = out["reference_port_one_at_a_time"]["decisions_per_s"]
        out["break_even_concurrency"] = next((int(c) for c in ("8", "64", "512", "4096")
                                              if conc[c] > per), None)
    # device side of the operating point between one request and the 64K headline: one routed
    # batch of n C3-shaped prompts, inputs resident, through the three-stream pipeline
    from b200_router.packing import pack_arrays
    from b200_router.trie import GpuHashTrie, RoutePipeline
    torch = cx.torch
    sweep = {}
    for nreq in (16, 256, 4096, 32768):
        trs = traces.SharedPrefixTrace(n_total=24 * nreq, batch=nreq, n_sys=64, sys_chars=SYS_CHARS,
                                       suffix_lo=SUFFIX[0], suffix_hi=SUFFIX[1], seed=SEED + 11)
        packs = [pack_arrays(*trs.batch_arrays(b), pin=False).to(cx.dev, non_blocking=False) for b in range(24)]
        pipe = RoutePipeline(GpuHashTrie(cap_log2=22, device=cx.dev), depth=3)
        pipe.reserve(nreq, max(p.max_chunks for p in packs))
        for b in range(8):
            pipe.route(packs[b], (1 << N_ENDPOINTS) - 1, SEED, b * nreq, inputs_ready=True)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for b in range(8, 24):
            pipe.route(packs[b], (1 << N_ENDPOINTS) - 1, SEED, b * nreq, inputs_ready=True)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 16
        sweep[str(nreq)] = {"us_per_batch": ms * 1e3, "decisions_per_s": nreq / (ms * 1e-3)}
        pipe.close()
        pipe.trie.close()
    out["device_batch_sweep"] = sweep
    out["config"] = {"workload": "C3-shaped prompts (2048-char shared system prompt + 256-1024-char "
                                 "suffix), 8 endpoints, Python str in, url out"}
    return out


def run_extras(cx, args, router, seq):
    ex = {}
    try:
        if cx.world == 1 and cx.rank == 0:
            ex["session"] = session_line(cx, args, K=10, W=3)
        # configs[0] (the reference's own CPU-runnable case: 10K multi-round-QA prompts, 4
        # endpoints), configs[4] (the larger replay), configs[3] (32K-char prompts, 64 endpoints,
        # shared-fraction sweep, exchange bytes and time at N > 1)
        if cx.world == 1:
            ex["c1"], seq = mrqa_line(cx, args, router, seq, users=1000, rounds=10, label="C1")
        ex["mrqa"], seq = mrqa_line(cx, args, router, seq)
        ex["c4"], seq = c4_line(cx, args, router, seq, fracs=(0.0, 0.5, 0.94), steps=3)
        if cx.world == 1 and cx.rank == 0:
            ex["latency"] = latency_line(cx, args)
    except Exception as exc:                               # extras never take the headline down
        import traceback
        ex["error"] = "%s: %s" % (type(exc).__name__, exc)
        traceback.print_exc(file=sys.stderr)
    return ex


def run_standalone(args):
    """--workload session | mrqa | c4 | latency on their own (the same functions as the extras)."""
    cx = Ctx()
    line = None
    if args.workload == "session":
        if cx.rank == 0:
            line = session_line(cx, args)
    elif args.workload == "latency":
        if cx.rank == 0:
            line = latency_line(cx, args)
    else:
        chunks = max(8192 * 256, 1 << 21)
        router = new_router(cx, max(22, min(28, 22 + int(np.ceil(np.log2(cx.world))) + 2)), 1 << 16, chunks)
        if args.workload == "mrqa":
            line, _ = mrqa_line(cx, args, router, 0)
        else:
            line, _ = c4_line(cx, args, router, 0, steps=max(2, min(args.steps, 4)))
    if cx.rank == 0 and line is not None:
        print(json.dumps(line))
    if cx.world > 1:
        cx.dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    elif a.workload == "prefix":
        run_b200(a)
    else:
        run_standalone(a)
