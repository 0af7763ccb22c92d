import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "vllm-production-stack_b200"))
import numpy as np, torch
from b200_router import traces
from b200_router.packing import pack_arrays
from b200_router.trie import GpuHashTrie, RoutePipeline
from oracle import coracle
tr = traces.SharedPrefixTrace(n_total=1 << 18, batch=1 << 16, n_sys=64, sys_chars=2048, suffix_lo=256, suffix_hi=1024, seed=7)
from b200_router import _lib
for cm in (1, 3):
  _lib.lib().b200r_set_claim_mode(cm)
  for mode in ("direct",):
    for cap in (22, 24):
          t = GpuHashTrie(cap_log2=cap)
          pipe = RoutePipeline(t)
          orc = coracle.TrieOracle()
          seq = 0
          for b in range(3):
              buf, starts, lens = tr.batch_arrays(b)
              if mode == "direct":
                  ep, depth, matched = t.route(pack_arrays(buf, starts, lens, pin=False), 0xFF, 99, seq)
              elif mode == "pipe_dev":
                  p = pack_arrays(buf, starts, lens, pin=False).to(t.device); torch.cuda.synchronize()
                  ep, depth, matched = pipe.route(p, 0xFF, 99, seq, inputs_ready=True)
              else:
                  ep, depth, matched = pipe.route(pack_arrays(buf, starts, lens, pin=True), 0xFF, 99, seq)
              oep, odepth, om = orc.route_batch(buf, starts, lens, 0xFF, 99, seq)
              seq += len(starts)
              e, d = ep.cpu().numpy(), depth.cpu().numpy()
              bad = np.nonzero((e != oep) | (d != odepth))[0]
              print(cm, mode, cap, b, "bad", len(bad), bad[:8], e[bad[:4]], oep[bad[:4]], d[bad[:4]], odepth[bad[:4]])
          st = t.status()
          print("  status", st, "oracle nodes", orc.num_nodes() - 1)
          t.close()
