set -x
ncu --set full --clock-control none --import-source on -k regex:frontend_coop_kernel --launch-skip 3 -c 1 -f -o gpurun_out/r1c_fe_f64 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --e2e-steps 1 --gl-batch 0 > gpurun_out/r1c_ncu_fe.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:frontend_coop_kernel --launch-skip 3 -c 1 -f -o gpurun_out/r1c_fe_f32 python bench.py --precision f32 --steps 1 --warmup 3 --no-cpu-baseline --e2e-steps 1 --gl-batch 0 > gpurun_out/r1c_ncu_fe32.log 2>&1
