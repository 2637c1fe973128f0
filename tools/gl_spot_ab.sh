# usage: bash tools/gl_spot_ab.sh N -- Griffin-Lim parity against the oracle on N of the 64 C3 spectrograms (100 classic / 32 momentum
# iterations) for the default build and every build/lib_gl_*.so: the distribution of a chaotic measure, not one draw
N=${1:-16}
for so in unast_b200/libunast_spectral.so build/lib_gl_*.so; do
  [ -f "$so" ] || continue
  UNAST_SPECTRAL_LIB=$PWD/$so python bench.py --steps 2 --warmup 3 --utts 300 --e2e-steps 1 --longform-seconds 0 --loader-batch 0 --io-utts 0 --gl-spot $N 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); p=d['parity_spot_check']; print('$so', 'gl', p['gl']['snr_db'], 'same S', p['gl'].get('same_S_snr_db'), 'fast', p['gl_fast']['snr_db'])"
done
