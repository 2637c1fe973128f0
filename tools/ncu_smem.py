"""Per-instruction shared-memory wavefronts of one kernel from an ncu report (needs --import-source on):
    python tools/ncu_smem.py REPORT.ncu-rep [frames] [top]
Prints the totals (wavefronts, excessive = bank conflicts) per frame and the instructions with the most excess."""
import csv, io, subprocess, sys
rep = sys.argv[1]
frames = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
lines = out.split("\n")
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))
def f(r, k):
    try: return float(r.get(k) or 0)
    except ValueError: return 0.0
tot = sum(f(r, "L1 Wavefronts Shared") for r in rows)
exc = sum(f(r, "L1 Wavefronts Shared Excessive") for r in rows)
inst = sum(f(r, "Instructions Executed") for r in rows)
print("instructions/frame %.1f  shared wavefronts/frame %.1f  excessive/frame %.1f (%.1f %%)" % (inst / frames, tot / frames, exc / frames, 100 * exc / max(tot, 1)))
rows.sort(key=lambda r: -f(r, "L1 Wavefronts Shared Excessive"))
for r in rows[:top]:
    print("%8.2f exc %8.2f wf %7.2f inst  %s" % (f(r, "L1 Wavefronts Shared Excessive") / frames, f(r, "L1 Wavefronts Shared") / frames,
                                              f(r, "Instructions Executed") / frames, r["Source"][:100]))
