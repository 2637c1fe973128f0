"""Per-region stall/instruction summary of an ncu source page: python tools/ncu_region.py rep lo hi [list]"""
import csv, io, subprocess, sys
path, lo, hi = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(out))); hdr, data = rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
S, X = col["# Samples"], col["Instructions Executed"]
stall = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
sub = data[lo:hi]
tot = sum(int(r[S]) for r in sub); allt = sum(int(r[S]) for r in data)
print("region [%d,%d): samples %d (%.1f%% of all), warp-instr executed %d" % (lo, hi, tot, 100.0 * tot / allt, sum(int(r[X]) for r in sub)))
agg = {h: sum(int(r[col[h]]) for r in sub) for h in stall}
print(", ".join("%s %.1f%%" % (h[6:], 100.0 * v / max(tot, 1)) for h, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
if len(sys.argv) > 4:
    ops = {}
    for r in sub:
        op = r[col["Source"]].strip().split()
        op = [o for o in op if not o.startswith("@")][0].split(".")[0]
        ops[op] = ops.get(op, 0) + int(r[X])
    print(sorted(ops.items(), key=lambda kv: -kv[1])[:25])
