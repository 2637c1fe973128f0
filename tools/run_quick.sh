# usage: bash tools/run_quick.sh TAG [extra bench args] -- GPU parity tests + bench without the CPU baseline
TAG=${1:-x}; shift
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
python bench.py --steps 5 --no-cpu-baseline --e2e-steps 1 "$@" > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc $?"
python -c "
import json;d=json.load(open('gpurun_out/${TAG}_bench.json'));print('FE ms',d['ms_per_step'],d['roofline']['kernel_ms'],d['parity_spot_check']);g=d.get('gl');print('GL ms',g and g['ms_per_step'],g and g['roofline']['frac'])"
