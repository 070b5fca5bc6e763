"""Shared-memory wavefronts per source line of one kernel (excessive = bank conflicts), from an ncu report
(development aid).   python tools/ncu_conflicts.py <report.ncu-rep> <cubin> <kernel substring> [launch index] [top]"""
import csv, io, re, subprocess, sys
from collections import defaultdict

rep, cubin, kname = sys.argv[1:4]
skip = sys.argv[4] if len(sys.argv) > 4 else "0"
top = int(sys.argv[5]) if len(sys.argv) > 5 else 25
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
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m:
        line_of[int(m.group(1), 16)] = (cur, m.group(2).split()[0] if m.group(2) else "")
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", skip, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
wi, ei = hdr.index("L1 Wavefronts Shared"), hdr.index("L1 Wavefronts Shared Excessive")
wave, exc = defaultdict(float), defaultdict(float)
base = None
for r in rows[hi + 1:]:
    if len(r) != len(hdr):
        continue
    a = int(r[0], 16)
    if base is None:
        base = a
    key = line_of.get(a - base, (None, ""))
    k = (key[0], key[1])
    wave[k] += float(r[wi] or 0)
    exc[k] += float(r[ei] or 0)
tw, te = sum(wave.values()), sum(exc.values())
print(f"shared wavefronts {tw:.3e}, excessive {te:.3e}")
for k, v in sorted(exc.items(), key=lambda kv: -kv[1])[:top]:
    print(f"{100 * v / max(te, 1):5.1f}% of excessive  {v:.3e} / {wave[k]:.3e} wavefronts  {k}")
