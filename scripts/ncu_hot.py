"""top stall locations of one kernel of an .ncu-rep (SASS view): ncu_hot.py <rep> <launch-skip> [n]"""
import csv, io, subprocess, sys
rep, skip = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", skip, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
print(rows[0][1][:100])
hdr = rows[1]; data = [r for r in rows[2:] if len(r) == len(hdr)]
si = hdr.index("# Samples"); ii = hdr.index("Instructions Executed")
tot = sum(int(r[si]) for r in data if r[si].isdigit())
print("total samples", tot, "instructions", len(data))
top = sorted([(int(r[si]), i, r[1].strip(), r[ii]) for i, r in enumerate(data) if r[si].isdigit()], reverse=True)[:n]
for s, i, src, ne in top:
    print(f"{s:7d} {100 * s / tot:5.1f}% idx{i:5d} exec {ne:>8s}  {src[:100]}")
