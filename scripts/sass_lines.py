"""Static instruction counts of the batch kernel by source region, from the in-tree library's SASS
(cuobjdump -xelf + nvdisasm --print-line-info; the library is built with -lineinfo). CPU only.
usage: python scripts/sass_lines.py [instantiation tag, default ILb0ELb0E]"""
import collections
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "ILb0ELb0E"
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "kubernetes_b200/lib/libksched_b200.so")], cwd=tmp,
                      stdout=subprocess.DEVNULL)
sass = subprocess.check_output(["nvdisasm", "--print-line-info", os.path.join(tmp, "kernels.sm_100a.cubin")], text=True)
cur_fn = cur_line = None
counts = collections.defaultdict(collections.Counter)
ops = collections.defaultdict(collections.Counter)
for ln in sass.split("\n"):
    m = re.match(r"\s*\.text\.(\S+):", ln)
    if m:
        cur_fn = m.group(1)
        continue
    m = re.search(r'//## File ".*kernels\.cu", line (\d+)', ln)
    if m:
        cur_line = int(m.group(1))
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)", ln)
    if m and cur_fn and cur_line:
        counts[cur_fn][cur_line] += 1
        ops[cur_fn][(cur_line, m.group(2).split(".")[0])] += 1
fn = [f for f in counts if tag in f][0]
c = counts[fn]
print(fn, "instructions:", sum(c.values()))
src = open(os.path.join(ROOT, "kubernetes_b200/csrc/engine/kernels.cu")).read().split("\n")


def find(txt):
    for i, l in enumerate(src):
        if txt in l:
            return i + 1
    raise KeyError(txt)


def rng(a, b):
    return sum(v for k, v in c.items() if a <= k < b)


regions = [
    ("make_view", "__device__ __forceinline__ PodView make_view", "__device__ __forceinline__ int pool_words"),
    ("unused_score", "__device__ __forceinline__ int64_t unused_score", "// 1/a for a in"),
    ("fast_rcp+lr_bra", "__device__ __forceinline__ double fast_rcp", "// The fork's per-node device selection"),
    ("select_devices*", "__device__ __forceinline__ uint32_t select_devices(", "__device__ __forceinline__ bool eval_expr("),
    ("eval_expr*", "__device__ __forceinline__ bool eval_expr(", "// PodMatchNodeSelector (predicates.go:691-761) and the NodeAffinityPriority Map"),
    ("load_row_common", "__device__ __forceinline__ void load_row_common", "// eval_expr for label_words == kLabFast"),
    ("eval_node", "__device__ __forceinline__ void eval_node", "// PrioritizeNodes' Reduce functions"),
    ("final_score", "__device__ __forceinline__ int32_t final_score", "__device__ __forceinline__ uint32_t ld_acquire("),
    ("eval_cached", "__device__ __forceinline__ void eval_cached", "// One warp: load row r"),
    ("select_merged", "__device__ __forceinline__ int select_merged", "// dynamic shared memory carve-up"),
    ("kernel: staging", "__global__ void __launch_bounds__(kThreads) ksb_schedule_batch_kernel", "// ---- phase 1: predicates + Map"),
    ("kernel: phase 1", "// ---- phase 1: predicates + Map", "// ---- phase 2: Reduce + weighted sum"),
    ("kernel: phase 2", "// ---- phase 2: Reduce + weighted sum", "// ---- phase 3: block 0 of rank 0 commits"),
    ("kernel: phase 3 set-up", "// ---- phase 3: block 0 of rank 0 commits", "// ================= one pod, exact (warp 0)"),
    ("kernel: phase 3 one-pod path", "// ================= one pod, exact (warp 0)", "// ================= a window of plain pods"),
    ("kernel: phase 3 picks", "// (1) picks, one pod per lane", "// (2) per pick: its row after the assume step"),
    ("kernel: phase 3 columns", "// (2) per pick: its row after the assume step", "// (3) how many of the window"),
    ("kernel: phase 3 finalize", "// (3) how many of the window", "// writer warp: table write-through"),
    ("kernel: writer", "// writer warp: table write-through", "      // publish how far the group got"),
]
for name, a, b in regions:
    try:
        print("%-32s lines %5d-%5d  %6d" % (name, find(a), find(b), rng(find(a), find(b))))
    except KeyError as e:
        print(name, "marker not found", e)
if len(sys.argv) > 2:  # opcode histogram of one region: name prefix
    for name, a, b in regions:
        if name.startswith(sys.argv[2]):
            h = collections.Counter()
            for (line, op), v in ops[fn].items():
                if find(a) <= line < find(b):
                    h[op] += v
            print(name, h.most_common(25))
