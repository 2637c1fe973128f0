# usage: bash tools/gl_ab.sh -- C3 Griffin-Lim timing only (no parity tests) for the default build and every build/lib_gl_*.so
for so in unast_b200/libunast_spectral.so build/lib_gl_*.so; do
  [ -f "$so" ] || continue
  UNAST_SPECTRAL_LIB=$PWD/$so python bench.py --steps 3 --warmup 3 --utts 300 --no-cpu-baseline --e2e-steps 1 --longform-seconds 0 --loader-batch 0 --io-utts 0 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); g=d['gl']; print('$so', 'gl ms', round(g['ms_per_step'],3), 'fast ms', round(g['fast']['ms_per_step'],3))"
done
