"""Per-source-line instruction shares of one kernel from an ncu report, for kernels whose code spans several files
(development aid).   python tools/ncu_lines2.py <report.ncu-rep> <cubin> <kernel substring> [launch index] [top]"""
import csv, io, re, subprocess, sys
from collections import defaultdict

rep, cubin, kname = sys.argv[1:4]
skip = sys.argv[4] if len(sys.argv) > 4 else "0"
top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
line_of, cur, infunc = {}, None, False
for ln in dis.splitlines():
    m = re.match(r"\s*\.text\.(\S+):", ln)
    if m:
        infunc = kname in m.group(1)
        continue
    if not infunc:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        f = m.group(1).split("/")[-1]
        if f in ("text.cu", "text_cells.cuh", "text_span.cuh", "text_packed.cuh"):
            cur = (f, int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m:
        line_of[int(m.group(1), 16)] = cur
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", skip, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ci, si, ti = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
ins, smp, thr = defaultdict(float), defaultdict(float), defaultdict(float)
base = None
for r in rows[hi + 1:]:
    if r and r[0] in ("Address", "Kernel Name"):
        break
    if len(r) != len(hdr):
        continue
    a = int(r[0], 16)
    if base is None:
        base = a
    key = line_of.get(a - base)
    ins[key] += float(r[ci]); smp[key] += float(r[si]); thr[key] += float(r[ti])
tot, tots = sum(ins.values()), sum(smp.values())
src = {}
for f in ("text.cu", "text_cells.cuh", "text_span.cuh", "text_packed.cuh"):
    try:
        src[f] = open("transmogrifai_b200/csrc/" + f).read().splitlines()
    except OSError:
        src[f] = []
print(f"total warp instructions {tot:.3e}, samples {tots:.0f}")
for key, v in sorted(ins.items(), key=(lambda kv: -smp[kv[0]]) if __import__("os").environ.get("BY_SAMPLES") else (lambda kv: -kv[1]))[:top]:
    text = src[key[0]][key[1] - 1].strip()[:90] if key and key[1] - 1 < len(src.get(key[0], [])) else ""
    print(f"{100*v/tot:5.1f}% ins {100*smp[key]/max(tots,1):5.1f}% smp  lanes {thr[key]/max(v,1):4.1f}  {key}  {text}")
if len(sys.argv) > 6:
    # sums over line ranges of text_cells.cuh:  name:lo-hi,...
    for spec in sys.argv[6].split(","):
        name, rng = spec.split(":")
        lo, hi = map(int, rng.split("-"))
        f = "text.cu" if name.startswith("T_") else "text_span.cuh" if name.startswith("S_") else "text_cells.cuh"
        v = sum(x for k, x in ins.items() if k and k[0] == f and lo <= k[1] <= hi)
        t = sum(x for k, x in thr.items() if k and k[0] == f and lo <= k[1] <= hi)
        print(f"{name:24s} {100*v/tot:5.1f}% ins   lanes {t/max(v,1):4.1f}")
    print("unattributed", 100 * ins.get(None, 0) / tot)
