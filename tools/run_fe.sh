# usage: bash tools/run_fe.sh TAG  -- GPU parity tests, f64 bench (front end only) and a full ncu capture of the front-end kernel
TAG=${1:-x}
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
python bench.py --steps 5 --no-cpu-baseline --e2e-steps 1 --gl-batch 0 > gpurun_out/${TAG}_bench_fe.json 2> gpurun_out/${TAG}_bench_fe.err; echo "bench rc $?"
python -c "
import json;d=json.load(open('gpurun_out/${TAG}_bench_fe.json'));print(d['ms_per_step'],d['roofline']['kernel_ms'],d['roofline']['kernel'],d['parity_spot_check'])"
ncu --set full --clock-control none --import-source on -k regex:frontend_ --launch-skip 3 -c 1 -f -o gpurun_out/${TAG}_fe_f64 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --e2e-steps 1 --gl-batch 0 > gpurun_out/${TAG}_ncu_fe.log 2>&1
