"""Basic-block view of one kernel from an ncu report: runs of SASS instructions with the same execution count, with
the source line of their first instruction (development aid).

    python tools/ncu_blocks.py <report.ncu-rep> <cubin> <kernel substring> <units, e.g. rows> [launch index]"""
import csv, io, re, subprocess, sys

rep, cubin, kname, units = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4])
skip = sys.argv[5] if len(sys.argv) > 5 else "0"
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
        if m.group(1).endswith(".cu"):
            cur = int(m.group(2))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m:
        line_of[int(m.group(1), 16)] = cur
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", skip, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ci, ti, si = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
data = []
for r in rows[hi + 1:]:
    if r and r[0] in ("Address", "Kernel Name"):
        break
    if len(r) == len(hdr):
        data.append((int(r[0], 16), float(r[ci] or 0), float(r[ti] or 0), float(r[si] or 0)))
base = data[0][0]
tot = sum(d[1] for d in data)
src = open(cubin.replace("/tmp/cub/", "/root/repo/transmogrifai_b200/csrc/").replace(".sm_100a.cubin", ".cu")).read().splitlines()
blocks, i = [], 0
while i < len(data):
    j = i
    while j + 1 < len(data) and abs(data[j + 1][1] - data[i][1]) <= 0.02 * max(data[i][1], 1):
        j += 1
    n = j - i + 1
    ins = sum(d[1] for d in data[i:j + 1])
    thr = sum(d[2] for d in data[i:j + 1])
    smp = sum(d[3] for d in data[i:j + 1])
    blocks.append((data[i][0] - base, n, data[i][1] / units, thr / max(ins, 1), ins / tot * 100, smp, line_of.get(data[i][0] - base)))
    i = j + 1
tsmp = sum(d[3] for d in data)
print(f"total {tot:.4g} warp instructions = {tot / units:.1f} per unit; {tsmp:.0f} stall samples")
for a, n, cnt, act, share, smp, ln in blocks:
    if share < 0.25 and smp / tsmp < 0.004:
        continue
    text = src[ln - 1].strip()[:70] if ln and ln <= len(src) else ""
    print(f"{a:05x} n={n:3d} x{cnt:8.1f} act {act:4.1f} ins {share:5.1f}% smp {smp / tsmp * 100:5.1f}%  {n * cnt:8.0f}/unit | {ln}: {text}")
