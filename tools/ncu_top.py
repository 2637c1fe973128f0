"""Top stall instructions of a `--set full --import-source on` capture (ncu --page source), grouped for quick reading.
    python tools/ncu_top.py gpurun_out/x.ncu-rep [N]"""
import csv, io, subprocess, sys
path = sys.argv[1]; N = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, data = rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
S, X = col["# Samples"], col["Instructions Executed"]
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[S]) for r in data)
print("total samples %d, SASS instructions %d, executed warp-instr %d" % (tot, len(data), sum(int(r[X]) for r in data)))
agg = {h: sum(int(r[col[h]]) for r in data) for h in stall_cols}
print("stall totals:", ", ".join("%s %.1f%%" % (h[6:], 100.0 * v / tot) for h, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
top = sorted(range(len(data)), key=lambda i: -int(data[i][S]))[:N]
for i in sorted(top):
    r = data[i]
    main = max(stall_cols, key=lambda h: int(r[col[h]]))
    print("%5d %7s %5.1f%% %-12s exec %-10s %s" % (i, r[S], 100.0 * int(r[S]) / tot, main[6:], r[X], r[col["Source"]].strip()[:80]))
