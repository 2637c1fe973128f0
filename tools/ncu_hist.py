"""Coarse histogram of an ncu source page: python tools/ncu_hist.py rep frames [bin]  -- per SASS-index bin: share of
stall samples, warp-instructions per frame, dominant stall reasons and opcodes."""
import csv, io, subprocess, sys
path, frames = sys.argv[1], float(sys.argv[2]); B = int(sys.argv[3]) if len(sys.argv) > 3 else 150
out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(out))); hdr, data = rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
S, X = col["# Samples"], col["Instructions Executed"]
stall = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[S]) for r in data)
print("total samples %d, executed warp-instr/frame %.1f" % (tot, sum(int(r[X]) for r in data) / frames))
for lo in range(0, len(data), B):
    sub = data[lo:lo + B]
    s = sum(int(r[S]) for r in sub); x = sum(int(r[X]) for r in sub)
    if s < tot * 0.005: continue
    agg = {h: sum(int(r[col[h]]) for r in sub) for h in stall}
    top = sorted(agg.items(), key=lambda kv: -kv[1])[:4]
    ops = {}
    for r in sub:
        op = [o for o in r[col["Source"]].split() if not o.startswith("@")][0].split(".")[0]
        ops[op] = ops.get(op, 0) + int(r[X])
    t3 = sorted(ops.items(), key=lambda kv: -kv[1])[:5]
    print("%5d-%5d samp %5.1f%% instr/frame %6.1f | %s | %s" % (lo, lo + B, 100 * s / tot, x / frames,
          ", ".join("%s %.0f%%" % (h[6:], 100 * v / max(s, 1)) for h, v in top), " ".join("%s:%d" % (k, v / frames) for k, v in t3)))
