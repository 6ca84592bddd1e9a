"""Join an `ncu --set full --import-source on` capture's per-instruction stall samples with the source lines of the
in-tree library (nvdisasm --print-line-info on the same build). Prints samples per source region and the top lines.
usage: python scripts/ncu_by_line.py REPORT.ncu-rep [kernel tag, default ILb0ELb0E] [top N]"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep = sys.argv[1]
tag = sys.argv[2] if len(sys.argv) > 2 else "ILb0ELb0E"
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.check_output(["ncu", "-i", rep, "--page", "source", "--csv"], text=True, stderr=subprocess.DEVNULL)
rows = list(csv.reader(io.StringIO(raw)))
hdr_i = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hdr_i]
col = {n: i for i, n in enumerate(hdr)}
inst = []
for r in rows[hdr_i + 1:]:
    if len(r) < len(hdr) or not r[0].startswith("0x"):
        continue
    inst.append((int(r[0], 16), r[1].strip(), int(r[col["# Samples"]] or 0), int(r[col["Instructions Executed"]] or 0), r))
base = inst[0][0]
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "kubernetes_b200/lib/libksched_b200.so")], cwd=tmp, stdout=subprocess.DEVNULL)
sass = subprocess.check_output(["nvdisasm", "--print-line-info", os.path.join(tmp, "kernels.sm_100a.cubin")], text=True)
line_of = {}
cur_fn = cur_line = None
for ln in sass.split("\n"):
    m = re.match(r"\s*\.text\.(\S+):", ln)
    if m:
        cur_fn = m.group(1)
        continue
    m = re.search(r'//## File ".*kernels\.cu", line (\d+)', ln)
    if m:
        cur_line = int(m.group(1))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+\S", ln)
    if m and cur_fn and tag in cur_fn:
        line_of[int(m.group(1), 16)] = cur_line
per_line = collections.Counter()
per_line_inst = collections.Counter()
stall_cols = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
per_line_stall = collections.defaultdict(collections.Counter)
miss = 0
for addr, txt, smp, ex, r in inst:
    ln = line_of.get(addr - base)
    if ln is None:
        miss += smp
        continue
    per_line[ln] += smp
    per_line_inst[ln] += ex
    for sc in stall_cols:
        v = int(r[col[sc]] or 0)
        if v:
            per_line_stall[ln][sc] += v
total = sum(per_line.values())
src = open(os.path.join(ROOT, "kubernetes_b200/csrc/engine/kernels.cu")).read().split("\n")
print("samples", total, "unmapped", miss, "warp instructions executed", sum(per_line_inst.values()))


def find(txt):
    for i, l in enumerate(src):
        if txt in l:
            return i + 1
    return None


marks = [("helpers (inlined everywhere)", 1), ("kernel: staging", find("__global__ void __launch_bounds__(kThreads) ksb_schedule_batch_kernel")),
         ("phase 1", find("// ---- phase 1: predicates + Map")), ("phase 2", find("// ---- phase 2: Reduce + weighted sum")),
         ("phase 3 set-up", find("// ---- phase 3: block 0 of rank 0 commits")),
         ("phase 3 one-pod path", find("// ================= one pod, exact (warp 0)")),
         ("phase 3 window", find("// ================= a window of plain pods")),
         ("phase 3 picks", find("// (1) picks, one pod per lane")), ("phase 3 columns", find("// (2) per pick: its row after the assume step")),
         ("phase 3 finalize", find("// (3) how many of the window")), ("writer", find("// writer warp: table write-through")),
         ("after phase 3", find("      // publish how far the group got")), ("end", len(src) + 1)]
for (n1, a), (_, b) in zip(marks, marks[1:]):
    s = sum(v for k, v in per_line.items() if a <= k < b)
    ix = sum(v for k, v in per_line_inst.items() if a <= k < b)
    print("%-30s lines %5d-%5d samples %8d (%5.1f%%)  warp-inst %12d" % (n1, a, b, s, 100.0 * s / max(1, total), ix))
print("top lines:")
for ln, s in per_line.most_common(topn):
    st = ", ".join("%s %d" % (k.replace("stall_", ""), v) for k, v in per_line_stall[ln].most_common(3))
    print("%6d %7d %5.1f%%  inst %10d | %s | %s" % (ln, s, 100.0 * s / total, per_line_inst[ln], st, src[ln - 1].strip()[:110]))
print("top lines by warp instructions executed:")
tot_inst = max(1, sum(per_line_inst.values()))
for ln, ix in per_line_inst.most_common(topn):
    print("%6d %12d %5.1f%%  samples %7d | %s" % (ln, ix, 100.0 * ix / tot_inst, per_line[ln], src[ln - 1].strip()[:110]))
