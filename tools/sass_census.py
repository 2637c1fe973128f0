"""Static SASS opcode census per kernel: cuobjdump -sass unast_b200/libunast_spectral.so | python tools/sass_census.py"""
import collections
import re
import sys

cur, cnt = None, collections.OrderedDict()
for line in sys.stdin:
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1).replace("_ZN5unast", "")
        cnt[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and cur:
        cnt[cur][m.group(1)] += 1
print("# SASS opcode census per shipped kernel (cuobjdump -sass unast_b200/libunast_spectral.so, sm_100a, static instruction counts)")
print("# Blackwell-specific: UBLKCP = cp.async.bulk (TMA), UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTCBAR = tcgen05.commit,")
print("# UTCATOMSWS = tcgen05.alloc / dealloc, FADD2 / FMUL2 / FFMA2 = packed f32x2, SYNCS = mbarrier; LDGSTS = cp.async.\n")
for k, c in cnt.items():
    print(k)
    print("    " + " ".join("%s:%d" % (o, n) for o, n in sorted(c.items())))
    print()
