# usage: bash tools/fe_ab.sh -- C2 front-end kernel time only for the default build, every build/lib_fe_*.so, and the default build again
for so in unast_b200/libunast_spectral.so build/lib_fe_*.so unast_b200/libunast_spectral.so; do
  UNAST_SPECTRAL_LIB=$PWD/$so python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 1 --gl-batch 0 --longform-seconds 0 --loader-batch 0 --io-utts 0 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$so', 'kernel_ms', round(d['roofline']['kernel_ms'],3))"
done
