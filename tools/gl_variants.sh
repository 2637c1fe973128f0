# usage: bash tools/gl_variants.sh -- Griffin-Lim parity tests + C3 timing for the default build and every build/lib_gl_*.so (A/B)
for so in unast_b200/libunast_spectral.so build/lib_gl_*.so; do
  [ -f "$so" ] || continue
  echo "== $so"
  UNAST_SPECTRAL_LIB=$PWD/$so timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "griffin or sharded or invert or irfft or rfft" 2>&1 | tail -1
  UNAST_SPECTRAL_LIB=$PWD/$so python bench.py --steps 3 --warmup 3 --utts 300 --no-cpu-baseline --e2e-steps 1 --longform-seconds 0 --loader-batch 0 --io-utts 0 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); g=d['gl']; print('gl ms', round(g['ms_per_step'],3), 'frac', round(g['roofline']['frac'],4), 'fast ms', round(g['fast']['ms_per_step'],3))"
done
