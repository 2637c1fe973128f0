# usage: bash tools/run_gl.sh TAG -- full ncu capture of one Griffin-Lim iteration launch (C3 batch)
TAG=${1:-x}
ncu --set full --clock-control none --import-source on -k regex:gl_ --launch-skip 150 -c 1 -f -o gpurun_out/${TAG}_gl python bench.py --steps 1 --warmup 3 --utts 500 --no-cpu-baseline --e2e-steps 1 > gpurun_out/${TAG}_ncu_gl.log 2>&1
echo rc $?
