# usage: bash tools/run_glm.sh TAG -- full ncu capture of one momentum-variant Griffin-Lim iteration launch (C3 batch)
TAG=${1:-x}
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:gl_pk_kernelILb1 --launch-skip 10 -c 1 -f -o gpurun_out/${TAG}_glm python bench.py --steps 1 --warmup 3 --utts 500 --no-cpu-baseline --e2e-steps 1 --gl-iters 4 --gl-steps 1 > gpurun_out/${TAG}_ncu_glm.log 2>&1
echo rc $?; tail -3 gpurun_out/${TAG}_ncu_glm.log
