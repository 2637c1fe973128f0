# usage: bash tools/run_scale.sh TAG N -- the driver's N-GPU launch of bench.py (one rank per GPU, no CPU baseline)
TAG=${1:-x}; N=${2:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_f64_${N}gpu.json 2> gpurun_out/${TAG}_bench_f64_${N}gpu.err
echo "rc $?"; tail -c 600 gpurun_out/${TAG}_bench_f64_${N}gpu.json | head -c 300; python -c "
import json;d=json.loads(open('gpurun_out/${TAG}_bench_f64_${N}gpu.json').read().strip().splitlines()[-1]);print(d['n_gpus'],d['value'],d['ms_per_step'],d['e2e']['value'],d['gl']['value'],d['gl']['ms_per_step'])"
