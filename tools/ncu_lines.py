"""Per-source-line instruction and stall-sample shares of one kernel from an ncu report (development aid).

    python tools/ncu_lines.py <report.ncu-rep> <cubin> <kernel substring> [launch index]

ncu's CSV source page lists SASS only; the line table of the cubin (nvdisasm -g) maps addresses back to text."""
import csv, io, re, subprocess, sys
from collections import defaultdict

rep, cubin, kname = sys.argv[1:4]
skip = sys.argv[4] if len(sys.argv) > 4 else "0"
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
        inl = "inlined" in ln
        cur = int(m.group(2)) if m.group(1).endswith("text.cu") or cur is None else cur
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
base = None
ins, smp, thr = defaultdict(float), defaultdict(float), defaultdict(float)
for r in rows[hi + 1:]:
    if r and r[0] in ("Address", "Kernel Name"):
        break
    if len(r) != len(hdr):
        continue
    a = int(r[0], 16)
    base = a if base is None else base
    ln = line_of.get(a - base)
    ins[ln] += float(r[ci] or 0)
    smp[ln] += float(r[si] or 0)
    thr[ln] += float(r[ti] or 0)
ti_, ts_ = sum(ins.values()), sum(smp.values())
src = open("/root/repo/transmogrifai_b200/csrc/text.cu").read().splitlines() if cubin.find("text") >= 0 else []
print(f"total warp instructions {ti_:.3g}, samples {ts_:.0f}")
for ln, v in sorted(ins.items(), key=lambda kv: -kv[1])[:50]:
    text = src[ln - 1].strip()[:95] if ln and ln <= len(src) else ""
    print(f"{v / ti_ * 100:5.1f}% ins {smp[ln] / ts_ * 100:5.1f}% smp  act {thr[ln] / max(v, 1):4.1f} | {ln}: {text}")
