# usage: bash tools/fe_variants.sh -- front-end parity tests + C2 kernel time for the default build and every build/lib_fe_*.so (A/B)
for so in unast_b200/libunast_spectral.so build/lib_fe_*.so; do
  [ -f "$so" ] || continue
  echo "== $so"
  UNAST_SPECTRAL_LIB=$PWD/$so timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "rfft or forward or ragged or tile or golden or int16 or short or real_speech or guard" 2>&1 | tail -1
  UNAST_SPECTRAL_LIB=$PWD/$so python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 1 --gl-batch 0 --longform-seconds 0 --loader-batch 0 --io-utts 0 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('kernel_ms', round(d['roofline']['kernel_ms'],3), 'mag', d['parity_spot_check']['max_abs_mag'], 'mel', d['parity_spot_check']['max_abs_mel'], 'bounds', d['parity_spot_check']['bounds_equal'])"
done
