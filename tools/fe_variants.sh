# usage: bash tools/fe_variants.sh -- times the front end of every build/lib_*.so variant on the C2 batch (A/B builds)
for so in build/lib_*.so; do
  UNAST_SPECTRAL_LIB=$PWD/$so python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 1 --gl-batch 0 --longform-seconds 0 --loader-batch 0 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$so', 'kernel_ms', round(d['roofline']['kernel_ms'],3), 'mag', d['parity_spot_check']['max_abs_mag'], 'mel', d['parity_spot_check']['max_abs_mel'], 'bounds', d['parity_spot_check']['bounds_equal'])"
done
