# usage: bash tools/run_glq.sh TAG -- GL parity tests + small-corpus bench (front end on 500 utterances), prints both GL objects
TAG=${1:-x}; shift
python -m pytest tests -m gpu -x -q -k "griffin or gl or momentum or vocoder" > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
python bench.py --steps 2 --utts 500 --no-cpu-baseline --e2e-steps 1 "$@" > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc $?"
python -c "
import json;d=json.load(open('gpurun_out/${TAG}_bench.json'));g=d.get('gl');print('GL ms',g['ms_per_step'],g['roofline']['frac']);print(g.get('fast'))"
