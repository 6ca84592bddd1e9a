"""ksched-b200: the kube-scheduler's per-pod node evaluation (genericScheduler.Schedule of
rotorliu/kubernetes) as sm_100a CUDA kernels behind a C-ABI.

Python is only the test / bench harness here: `lib` loads the in-tree shared libraries,
`workloads` generates the seeded synthetic clusters of BASELINE.json. There is no CPU
fallback: loading fails loudly when the CUDA library is missing.
"""
from . import lib, workloads  # noqa: F401
