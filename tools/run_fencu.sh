# usage: bash tools/run_fencu.sh TAG -- full ncu capture of the float64 front-end launch of a C2 bench step
TAG=${1:-x}
ncu --set full --clock-control none --import-source on -k regex:frontend_ --launch-skip 3 -c 1 -f -o gpurun_out/${TAG}_fe_f64 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --e2e-steps 1 --gl-batch 0 --loader-batch 0 --longform-seconds 0 > gpurun_out/${TAG}_ncu_fe.log 2>&1
echo rc $?
