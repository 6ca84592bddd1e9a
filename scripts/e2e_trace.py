"""Host-side timing breakdown of ksb_schedule_batch on a BASELINE.json config (KSB_TRACE=1 makes the engine print
its marks on stderr): python scripts/e2e_trace.py [CFG] [RUNS]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["KSB_TRACE"] = "1"
from kubernetes_b200 import lib, workloads  # noqa: E402

cfg = int(sys.argv[1]) if len(sys.argv) > 1 else 4
runs = int(sys.argv[2]) if len(sys.argv) > 2 else 3
nodes, pods = {1: (100, 1000), 2: (1000, 10000), 3: (5000, 30000), 4: (50000, 100000), 5: (5000, 20000)}[cfg]
sess = lib.Session(cluster=workloads.generate(cfg, nodes, pods))
for i in range(runs):
    out = sess.run(results=False)
    print({k: out[k] for k in ("e2e_ms", "kernel_ms", "encode_ms", "decode_ms", "post_ms", "sync_ms", "load_ms") if k in out}, flush=True)
sess.close()
