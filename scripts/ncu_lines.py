#!/usr/bin/env python
"""Per-CUDA-source-line instruction / sample shares for one kernel of an .ncu-rep, by joining the SASS page of the
report with the line table of the cubin inside the built .so (same build!).
usage: scripts/ncu_lines.py <report.ncu-rep> <kernel-name-regex> <mangled-function-substring> <cubin stem> [top] [section]
(section: which of the matching launches, 0-based)"""
import collections, csv, os, re, subprocess, sys, tempfile

rep, kre, fsub, stem = sys.argv[1:5]
top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(root, "bwa-mem-gpu_b200", "libbwamem_b200.so")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
cub = [f for f in os.listdir(tmp) if f.startswith(stem + ".")][0]
sass = subprocess.run(["nvdisasm", "-g", os.path.join(tmp, cub)], capture_output=True, text=True).stdout.split("\n")
seq, cur, infn = [], None, False
for ln in sass:
    if ln.startswith(".text."):
        infn = fsub in ln
        continue
    if ln.startswith("//-----"):
        infn = False
    if not infn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m:
        seq.append((cur, m.group(2)))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kre],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.split("\n")))
h = rows[1]
I = h.index
sect = int(sys.argv[6]) if len(sys.argv) > 6 else 0
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
print(f"{len(starts) - 1} sections in the report page; using section {sect}")
data = [r for r in rows[starts[sect] + 2: starts[sect + 1]] if len(r) > 10]
if len(data) != len(seq):
    print(f"warning: {len(data)} SASS rows in report vs {len(seq)} in cubin (different build?)")
agg = collections.defaultdict(lambda: [0, 0, 0])
tot = ts = 0
for r, (cur, txt) in zip(data, seq):
    inst, thr, s = int(r[I("Instructions Executed")]), int(r[I("Thread Instructions Executed")]), int(r[I("# Samples")])
    a = agg[cur]
    a[0] += inst; a[1] += thr; a[2] += s
    tot += inst; ts += s
src = {}
print(f"total warp instructions {tot}, samples {ts}, lanes/inst {sum(a[1] for a in agg.values()) / tot:.1f}")
for cur, (inst, thr, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    f, l = cur if cur else ("?", 0)
    if f not in src:
        p = os.path.join(root, "bwa-mem-gpu_b200", "csrc", f)
        src[f] = open(p).read().split("\n") if os.path.exists(p) else None
    text = src[f][l - 1].strip()[:105] if src[f] and l <= len(src[f]) else ""
    print(f"{inst / tot * 100:5.1f}% lanes {thr / max(inst, 1):5.1f} samp {s / max(ts, 1) * 100:5.1f}% {f}:{l} {text}")
