# usage: bash tools/run_prof.sh TAG [fe|gl|both] -- plain run first (exit code checked), then one full ncu capture per hot kernel
TAG=${1:-x}; WHAT=${2:-both}
set -o pipefail
if [ "$WHAT" = fe ] || [ "$WHAT" = both ]; then
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --e2e-steps 1 --gl-batch 0 --longform-seconds 0 --loader-batch 0 > gpurun_out/${TAG}_plain_fe.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:frontend_ --launch-skip 3 -c 1 -f -o gpurun_out/${TAG}_fe python bench.py --steps 1 --warmup 3 --no-cpu-baseline --e2e-steps 1 --gl-batch 0 --longform-seconds 0 --loader-batch 0 > gpurun_out/${TAG}_ncu_fe.log 2>&1
echo "fe rc $?"
fi
if [ "$WHAT" = gl ] || [ "$WHAT" = both ]; then
python bench.py --steps 1 --warmup 3 --utts 500 --no-cpu-baseline --e2e-steps 1 --longform-seconds 0 --loader-batch 0 > gpurun_out/${TAG}_plain_gl.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gl_persist --launch-skip 1 -c 1 -f -o gpurun_out/${TAG}_gl python bench.py --steps 1 --warmup 3 --utts 500 --no-cpu-baseline --e2e-steps 1 --longform-seconds 0 --loader-batch 0 > gpurun_out/${TAG}_ncu_gl.log 2>&1
echo "gl rc $?"
fi
