#!/usr/bin/env python
"""bench.py -- prefix-route decisions/s (and prompt GB/s vs the HBM roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Workload (BASELINE.json configs[2], SURVEY.md 8d "C3"): prompts = one of 64
shared 2048-char system prompts + a unique 256..1024-char suffix, 8 endpoints,
batches of 65536 requests, one endpoint dropped for two batches (availability
churn).  A step = one batch through the whole routing path (hash, lookup,
tie-break, insert; results equal to the reference's sequential order).  Every
step routes a batch the table has not seen; batches (~177 MB of prompt bytes)
are larger than L2 (126 MB) and distinct, so no flush is needed between steps.

  value      whole-job decisions/s, batches resident in HBM, CUDA events
  e2e        same through the public Python API from pinned HOST buffers:
             H2D of the packed batch + route + D2H of the decisions, per step
  roofline   k_hash_chunks (the dominant kernel): prompt bytes per launch /
             its CUDA-event duration (second pass with per-kernel events)
  cpu_baseline  oracle/prefix_ref.py (Python restatement of the reference,
             asyncio structure included) on a bounded sample, 1 core

--impl reference times that restatement alone (the reference is pure Python
and is not installable on the GPU box).  Under torchrun every
rank runs the b200 arm on its shard; for --impl reference only rank 0 works.
"""
import argparse
import json
import os
import subprocess
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


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="prefix", choices=["prefix", "session", "mrqa"],
                    help="prefix = BASELINE metric (default); session = configs[1], 1M session ids; "
                         "mrqa = configs[4], multi-round-QA trace replay")
    ap.add_argument("--batch", type=int, default=BATCH, help=argparse.SUPPRESS)
    ap.add_argument("--cpu-sample", type=int, default=BATCH, help=argparse.SUPPRESS)
    ap.add_argument("--no-cpu", action="store_true", help=argparse.SUPPRESS)
    return ap.parse_args()


def workload_config(args, n_gpus):
    return {"workload": "C3: 64 shared 2048-char system prompts + unique 256-1024-char suffix, "
                        "%d endpoints, batch %d per GPU, endpoint 3 unavailable in batches 5-6"
                        % (N_ENDPOINTS, args.batch),
            "batch_per_gpu": args.batch, "global_batch": args.batch * n_gpus,
            "chunk_chars": 128, "endpoints": N_ENDPOINTS,
            "l2": "inputs larger than L2: each step reads a distinct ~%d MB batch"
                  % (args.batch * 2700 // 1000000),
            "tie_break": "canonical (endpoint-index order, counter-based draw)",
            "multi_gpu": None if n_gpus == 1 else
            {"": "table partitioned by first-chunk hash, NVLink peer-memory exchange",
             "replicated": "table replicated, NVLink peer-memory exchange of inserted nodes",
             "nccl": "table replicated, NCCL all-gather of inserted nodes"}[
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
        while not event.query() and len(self.samples) < limit:
            time.sleep(0.0003)
            if not event.query():
                self.sample()

    def result(self):
        out = {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
               "samples": len(self.samples)}
        if self.samples:
            out["sm_mhz"] = float(np.median(self.samples))
        return out


# ------------------------------------------------------- CPU (oracle) legs
def prompts_of(buf, starts, lens, count):
    out = []
    for i in range(min(count, len(starts))):
        s = int(starts[i])
        out.append(buf[s:s + int(lens[i])].tobytes().decode("ascii"))
    return out


def cpu_port_rate(prompts, avail, seed=SEED):
    """Python restatement of the reference router, one asyncio loop, 1 core.
    Returns (decisions/s, decisions as endpoint indices)."""
    from oracle import prefix_ref
    from oracle.chooser import ChoiceShim
    from b200_router import traces
    urls = traces.endpoint_urls(N_ENDPOINTS)
    index_of = {u: i for i, u in enumerate(urls)}
    router = prefix_ref.PrefixRouterRef(ChoiceShim(index_of, seed))
    eps = [urls[i] for i in range(N_ENDPOINTS) if (avail >> i) & 1]
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
        buf, starts, lens = tr.batch_arrays(b) if args.batch <= 8192 else \
            _first_requests(tr, b, sample)
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


def _first_requests(tr, b, count):
    """First `count` requests of batch b without materialising the whole batch."""
    from b200_router import traces
    small = traces.SharedPrefixTrace(n_total=tr.n_total, batch=tr.batch, n_sys=tr.n_sys,
                                     sys_chars=tr.sys_chars, suffix_lo=tr.suffix_lo,
                                     suffix_hi=tr.suffix_hi, seed=tr.seed)
    buf, starts, lens = small.batch_arrays(b)
    return buf, starts[:count], lens[:count]


# ------------------------------------------------------------- B200 arm
def run_b200(args):
    import torch
    import torch.distributed as dist
    from b200_router import _lib
    from b200_router.packing import pack_arrays
    from b200_router.trie import GpuHashTrie

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the B200 path has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    W, K = args.warmup, args.steps
    nb = W + K

    # ---- synthetic batches (host, pinned), this rank's shard of every global batch
    tr = make_trace(args, nb * world)
    host = []
    for b in range(nb):
        buf, starts, lens = tr.batch_arrays(b * world + rank)
        host.append(pack_arrays(buf, starts, lens, pin=True))
    max_chunks = max(p.max_chunks for p in host)

    nodes_est = nb * world * args.batch * 8
    cap_log2 = max(16, min(28, int(np.ceil(np.log2(nodes_est * 2.2)))))

    def new_router():
        if world == 1:
            # public single-GPU API: table + two-stream batch pipeline (hash stage of batch
            # b+1 overlaps the route stage of batch b; results stay sequential)
            from b200_router.trie import RoutePipeline
            pipe = RoutePipeline(GpuHashTrie(cap_log2=cap_log2, device=dev),
                                 depth=int(os.environ.get("BENCH_PIPE_DEPTH", "2")))
            pipe.reserve(args.batch, max_chunks)
            return pipe
        # multi-GPU: table partitioned by first-chunk hash, chunk hashes and results shipped
        # over NVLink peer memory by the library's own kernels (the variant that scales).
        # BENCH_EXCHANGE=replicated / nccl select the replicated-table variants (peer-memory
        # or NCCL all-gather exchange of inserted nodes) for comparison.
        from b200_router.distributed import PartitionedPrefixRouter, ShardedPrefixRouter
        cls = {"nccl": ShardedPrefixRouter, "replicated": ShardedPrefixRouter}.get(
            os.environ.get("BENCH_EXCHANGE", ""), PartitionedPrefixRouter)
        return cls(cap_log2=cap_log2, device=dev, max_requests=args.batch, max_chunks=max_chunks)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    out = (torch.empty(args.batch, dtype=torch.int32, device=dev),
           torch.empty(args.batch, dtype=torch.int32, device=dev),
           torch.empty(args.batch, dtype=torch.int64, device=dev))

    def route_resident(router, b, seq):
        if world == 1:
            return router.route(resident[b], avail_mask(b), SEED, seq, inputs_ready=True)
        if hasattr(router, "send_stream"):                       # partitioned: pipelined
            return router.route(resident[b], avail_mask(b), SEED, seq, out=out, inputs_ready=True)
        return router.route(resident[b], avail_mask(b), SEED, seq, out=out)

    # ---- pass 1: device-resident batches, whole timed region bracketed by events
    resident = [p.to(dev, non_blocking=False) for p in host]
    router = new_router()
    # NVML is initialised here, well before the barrier that opens the timed region: its
    # start-up takes milliseconds on the host and every rank waits for the slowest one
    sampler = ClockSampler(local, enabled=(rank == 0))
    sampler.sample()
    sampler.samples.clear()
    seq = 0
    for b in range(W):
        route_resident(router, b, seq)
        seq += args.batch * world
    barrier()
    launches0 = _lib.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    th0 = time.perf_counter()
    for b in range(W, nb):
        route_resident(router, b, seq)
        seq += args.batch * world
    host_submit_ms = (time.perf_counter() - th0) * 1e3 / K
    ev1.record()
    if rank == 0:
        sampler.sample_until(ev1)        # GPU is still inside the timed region here
    barrier()
    clocks = sampler.result()
    launches = _lib.launch_count() - launches0
    ms = ev0.elapsed_time(ev1)
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    st = router.check_status(count_nodes=True)
    n_dec = K * args.batch * world
    payload = sum(p.payload_bytes for p in host[W:])
    if world > 1:
        t = torch.tensor([payload], device=dev, dtype=torch.int64)
        dist.all_reduce(t)
        payload = int(t.item())
    value = n_dec / (ms * 1e-3)

    # ---- pass 2: the same W+K batches on a fresh table, with per-kernel CUDA events
    # around every launch of the K timed steps (roofline of k_hash_chunks, kernel shares).
    # Single stream here (no batch overlap), so that every kernel is timed on its own and
    # its share of the step is comparable with the serialised ncu launch list.
    del router
    router = new_router()
    serial = router.trie if world == 1 else router
    seq = 0
    for b in range(W):
        if world == 1:
            serial.route(resident[b], avail_mask(b), SEED, seq, out=out)
        else:
            route_resident(router, b, seq)
        seq += args.batch * world
    barrier()
    _lib.prof_enable(True)
    verbose = os.environ.get("BENCH_VERBOSE") == "1"
    prof = None
    for b in range(W, nb):
        if world == 1:
            serial.route(resident[b], avail_mask(b), SEED, seq, out=out)
        else:
            route_resident(router, b, seq)
        seq += args.batch * world
        if verbose:                      # per-step kernel times (synchronises every step)
            one = _lib.prof_read()
            print("step %d avail=%x " % (b, avail_mask(b)) +
                  " ".join("%s=%.1fus" % (k, v[0] * 1e3) for k, v in one.items() if v[1]),
                  file=sys.stderr)
            prof = one if prof is None else {k: (prof[k][0] + one[k][0], prof[k][1] + one[k][1])
                                             for k in one}
    torch.cuda.synchronize()
    if not verbose:
        prof = _lib.prof_read()
    _lib.prof_enable(False)
    my_payload = sum(p.payload_bytes for p in host[W:])
    hash_ms, hash_n = prof["hash"]
    peaks, peak_src = load_peaks()
    achieved = my_payload / (hash_ms * 1e-3) / 1e9 if hash_ms > 0 else 0.0
    prof_total = sum(v[0] for v in prof.values())
    del router
    resident = None
    torch.cuda.empty_cache()

    # ---- pass 3: end to end from pinned host buffers through the public API.  Every step
    # copies its packed batch host->device and its decisions device->host inside the timed
    # region; the host reads step b's decisions after submitting step b+1 (the pipeline
    # overlaps the copy of the next batch with the routing of the current one).
    router = new_router()
    seq = 0
    res_host = [torch.empty(args.batch, dtype=torch.int32).pin_memory() for _ in range(3)]
    res_ev = [torch.cuda.Event() for _ in range(3)]

    def submit(b, seq, j):
        if world == 1:
            ep, _, _ = router.route(host[b], avail_mask(b), SEED, seq)
        elif hasattr(router, "send_stream"):      # host inputs are copied on the send stream
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
    barrier()
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
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    h2d = int(np.mean([p.h2d_bytes() for p in host[W:]]))
    e2e = {"value": n_dec / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": args.batch * 4, "ms_per_step": e2e_s / K * 1e3}

    # ---- the same from Python str objects (what a caller of route_batch holds): native packer
    #      (libb200pack.so) into a pinned arena + H2D + route + D2H, pipelined.  Host-bound:
    #      reported next to e2e, not instead of it.
    e2e_strings = None
    if rank == 0 and world == 1 and not args.no_cpu:
        from b200_router.packing import PackArena, PackPipeline
        ns = min(4, nb)
        strs = [prompts_of(*tr.batch_arrays(b), args.batch) for b in range(ns)]
        pp = PackPipeline(arena=PackArena(slots=5), workers=2)     # packs 2 batches ahead
        for rep in range(3):                      # reps 0-1 touch every arena slot (pinned allocation)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            prev = None
            for k in range(min(2, ns)):
                pp.submit(strs[k])
            for k in range(ns):
                pk = pp.next()
                if k + 2 < ns:
                    pp.submit(strs[k + 2])
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
                       "note": "input = list of Python str per step; native packer on 2 worker threads "
                               "(copy phase without the GIL, %d cpus) + H2D + route + D2H" % os.cpu_count()}
        del strs

    # ---- CPU baseline (rank 0, N=1 only) + parity of the same sample on a fresh table
    cpu = None
    parity = None
    if rank == 0 and world == 1 and not args.no_cpu:
        b = W
        cnt = min(args.cpu_sample, args.batch)
        buf, starts, lens = tr.batch_arrays(b)
        prompts = prompts_of(buf, starts, lens, cnt)
        rate, want = cpu_port_rate(prompts, avail_mask(b))
        router = None
        torch.cuda.empty_cache()
        t2 = GpuHashTrie(cap_log2=max(16, int(np.ceil(np.log2(cnt * 24 * 2.2)))), device=dev)
        ep, _, _ = t2.route(pack_arrays(buf, starts[:cnt], lens[:cnt], pin=False),
                            avail_mask(b), SEED, 0)
        parity = bool(ep.cpu().tolist() == want)
        t2.close()
        cpu = {"value": rate, "unit": UNIT, "cores": 1, "kind": "port",
               "sample": "first %d requests of timed batch 0 on an empty table; Python "
                         "restatement of the reference (oracle/prefix_ref.py), one asyncio loop, "
                         "host has %d cpus" % (cnt, os.cpu_count())}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K,
                "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u64", "data": "synthetic",
                "config": workload_config(args, world),
                "prompt_gbs": payload / (ms * 1e-3) / 1e9,
                "prompt_frac_of_hbm": payload / (ms * 1e-3) / 1e9 / peaks["hbm_gbs"],
                "clocks": clocks, "host_submit_ms_per_step": host_submit_ms, "e2e": e2e, "e2e_strings": e2e_strings,
                "gpu_launches": int(launches),
                "roofline": {"bound": "hbm", "kernel": "k_hash_chunks", "achieved": achieved,
                             "peak": peaks["hbm_gbs"], "unit": "GB/s",
                             "frac": achieved / peaks["hbm_gbs"],
                             "traffic": load_ncu_traffic("k_hash_chunks"),
                             "peak_source": peak_src, "launches": hash_n,
                             "ms_per_launch": hash_ms / max(hash_n, 1),
                             "share_of_step": hash_ms / prof_total if prof_total else None,
                             "kernel_ms_per_step": {k: v[0] / K for k, v in prof.items() if v[1]}},
                "cpu_baseline": cpu, "parity_sample_ok": parity,
                "table": {"nodes": st["num_nodes"], "capacity": st["capacity"]}}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def load_ncu_traffic(kernel):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return json.load(f)[kernel]["dram_bytes_per_launch"]
    except Exception:
        return None


def run_session(args):
    """configs[1]: 1M synthetic session ids over 16 endpoints, batched ring lookup on 1 GPU.
    ALU-bound (one MD5 block per key); the HBM fraction is informational."""
    import torch
    from b200_router import traces
    from b200_router.packing import pack_arrays
    from b200_router.ring import GpuHashRing
    n = 1 << 20
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    nodes = traces.endpoint_urls(16)
    ring = GpuHashRing(nodes)
    buf, starts, lens = traces.session_ids_packed(n, seed=0)
    host = pack_arrays(buf, starts, lens, pin=True)
    res = host.to(dev, non_blocking=False)
    W, K = args.warmup, args.steps
    for _ in range(W):
        out = ring.lookup_packed(res)
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(K):
        out = ring.lookup_packed(res)
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    t0 = time.perf_counter()
    for _ in range(K):
        got = ring.lookup_packed(host.to(dev, non_blocking=True)).cpu()
    e2e_s = time.perf_counter() - t0
    cpu = None
    if not args.no_cpu:
        from oracle import ring_ref
        ref = ring_ref.RingRef(nodes)
        ids = traces.session_ids_uuid(200000, 0)
        t0 = time.perf_counter()
        want = [ref.get_node(k) for k in ids]
        dt = time.perf_counter() - t0
        cpu = {"value": len(ids) / dt, "unit": "lookups/s", "cores": 1, "kind": "port",
               "sample": "200000 uuid ids through oracle/ring_ref.py (restated uhashring ring)"}
        assert ring.get_node_batch(ids[:5000]) == want[:5000]
    peaks, src = load_peaks()
    algo = n * (36 + 8 + 4 + 4)
    print(json.dumps({
        "metric": "session_route_lookups_per_sec", "value": n * K / (ms * 1e-3), "unit": "lookups/s",
        "n_gpus": 1, "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": "C2: 1M uuid-shaped session ids, 16 endpoints, ring of 2560 points"},
        "e2e": {"value": n * K / e2e_s, "unit": "lookups/s", "h2d_bytes_per_step": host.h2d_bytes(),
                "d2h_bytes_per_step": n * 4},
        "roofline": {"bound": "hbm", "kernel": "k_ring_lookup", "achieved": algo / (ms / K * 1e-3) / 1e9,
                     "peak": peaks["hbm_gbs"], "unit": "GB/s",
                     "frac": algo / (ms / K * 1e-3) / 1e9 / peaks["hbm_gbs"], "traffic": None,
                     "note": "ALU-bound by construction: one 64-step MD5 block per 36-byte key"},
        "md5_blocks_per_sec": n * K / (ms * 1e-3), "cpu_baseline": cpu, "gpu_launches": K}))


def run_mrqa(args):
    """configs[4]: replay of a benchmarks/multi-round-qa style trace (users x rounds, every
    round carries the whole chat history) through the GPU router, one batch per round, vs
    the Python restatement of the reference on a sample of the same trace (1 GPU)."""
    import torch
    from b200_router import traces
    from b200_router.packing import pack_prompts
    from b200_router.trie import GpuHashTrie, RoutePipeline
    users, rounds = 5000, 10
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    prompts, _ = traces.multi_round_qa(users=users, rounds=rounds, sys_len=1000, hist_len=200,
                                       answer_len=100)
    host = [pack_prompts(prompts[r * users:(r + 1) * users]) for r in range(rounds)]
    resident = [p.to(dev, non_blocking=False) for p in host]
    payload = sum(p.payload_bytes for p in host)
    avail = (1 << 4) - 1

    def run(pipe, batches, ready):
        seq = 0
        for p in batches:
            out = pipe.route(p, avail, SEED, seq, inputs_ready=ready)
            seq += p.n
        return out
    for _ in range(max(args.warmup, 1)):                # warm-up on throw-away tables
        t = GpuHashTrie(cap_log2=22, device=dev)
        run(RoutePipeline(t), resident, True)
        torch.cuda.synchronize()
        t.close()
    t = GpuHashTrie(cap_log2=22, device=dev)
    pipe = RoutePipeline(t)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    run(pipe, resident, True)
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    nodes = t.check_status(count_nodes=True)["num_nodes"]
    t.close()
    t = GpuHashTrie(cap_log2=22, device=dev)
    pipe = RoutePipeline(t)
    t0 = time.perf_counter()
    last = run(pipe, host, False)
    got = last[0].cpu()
    e2e_s = time.perf_counter() - t0
    t.close()
    cpu = None
    if not args.no_cpu:
        sample = prompts[:users * 2]                    # the first two rounds
        rate, want = cpu_port_rate_generic(sample, 4)
        t = GpuHashTrie(cap_log2=20, device=dev)
        ep0, _, _ = t.route(host[0].to(dev), avail, SEED, 0)
        ep1, _, _ = t.route(host[1].to(dev), avail, SEED, users)
        parity = (ep0.cpu().tolist() + ep1.cpu().tolist()) == want
        t.close()
        cpu = {"value": rate, "unit": UNIT, "cores": 1, "kind": "port", "parity_sample_ok": parity,
               "sample": "first 2 rounds (%d requests) through oracle/prefix_ref.py" % len(sample)}
    peaks, _ = load_peaks()
    n = users * rounds
    print(json.dumps({
        "metric": METRIC, "value": n / (ms * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": rounds,
        "warmup": max(args.warmup, 1), "ms_per_step": ms / rounds, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": "C5: multi-round-QA replay, %d users x %d rounds, 4 endpoints, one "
                               "batch per round, prompts %.1f KB on average" % (users, rounds,
                                                                                payload / n / 1e3)},
        "prompt_gbs": payload / (ms * 1e-3) / 1e9,
        "prompt_frac_of_hbm": payload / (ms * 1e-3) / 1e9 / peaks["hbm_gbs"],
        "e2e": {"value": n / e2e_s, "unit": UNIT, "h2d_bytes_per_step": sum(p.h2d_bytes() for p in host) // rounds,
                "d2h_bytes_per_step": users * 4},
        "cpu_baseline": cpu, "table": {"nodes": nodes}, "gpu_launches": int(_lib_launches())}))


def _lib_launches():
    from b200_router import _lib
    return _lib.launch_count()


def cpu_port_rate_generic(prompts, n_endpoints, seed=SEED):
    from oracle import prefix_ref
    from oracle.chooser import ChoiceShim
    from b200_router import traces
    urls = traces.endpoint_urls(n_endpoints)
    index_of = {u: i for i, u in enumerate(urls)}
    router = prefix_ref.PrefixRouterRef(ChoiceShim(index_of, seed))
    t0 = time.perf_counter()
    got = router.route_many(urls, prompts)
    dt = time.perf_counter() - t0
    return len(prompts) / dt, [index_of[u] for u in got]


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return json.load(f), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md)"


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    elif a.workload == "session":
        run_session(a)
    elif a.workload == "mrqa":
        run_mrqa(a)
    else:
        run_b200(a)
