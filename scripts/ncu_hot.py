"""Hot-spot summary of an ncu report: `python scripts/ncu_hot.py report.ncu-rep [launch-index]`.
Aggregates executed warp instructions by opcode and by straight-line block (runs of equal
execution count) from the SASS source page."""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]
k = sys.argv[2] if len(sys.argv) > 2 else "0"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass",
                      "--launch-skip", k, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
print(rows[0][:2])
hdr = rows[1]
iS, iE, iSm = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
data = []
for r in rows[2:]:
    try:
        data.append((int(r[iE]), int(r[iSm]), r[iS].strip()))
    except Exception:
        pass
tot = sum(d[0] for d in data)
def opc(src):
    t = src.split()
    o = t[1] if t[0].startswith("@") else t[0]
    return o.split(".")[0]
byop = collections.Counter()
for n, s, src in data:
    byop[opc(src)] += n
print("total warp instr", tot, "static", len(data))
print(" ".join(f"{k}:{v / tot:.3f}" for k, v in byop.most_common(16)))
blocks, cur = [], None
for i, (n, s, src) in enumerate(data):
    if cur is None or n != cur[0]:
        cur = [n, 0, 0, i, collections.Counter()]
        blocks.append(cur)
    cur[1] += 1; cur[2] += s; cur[4][opc(src)] += 1
blocks.sort(key=lambda b: -b[0] * b[1])
for b in blocks[:16]:
    print(f"@{b[3]:5d} len {b[1]:4d} exec {b[0]:10d} share {b[0] * b[1] / tot:.3f} samples {b[2]:7d}",
          dict(b[4].most_common(6)))
if len(sys.argv) > 3:
    a = int(sys.argv[3]); n = int(sys.argv[4]) if len(sys.argv) > 4 else 60
    for i in range(a, a + n):
        print(i, data[i][0], data[i][2])
