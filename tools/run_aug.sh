# usage: bash tools/run_aug.sh TAG -- augment / loader parity tests + small-corpus bench, prints the loader object
TAG=${1:-x}; shift
python -m pytest tests -m gpu -x -q -k "augment or specaugment or noise_fn or loader" > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/${TAG}_pytest.log
tail -15 gpurun_out/${TAG}_pytest.log
python bench.py --steps 2 --utts 500 --no-cpu-baseline --e2e-steps 1 --gl-batch 0 "$@" > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc $?"; tail -5 gpurun_out/${TAG}_bench.err
python -c "
import json;d=json.load(open('gpurun_out/${TAG}_bench.json'));print(json.dumps(d.get('loader'),indent=1))"
