set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r1c_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/r1c_pytest.log
python bench.py > gpurun_out/r1c_bench_f64.json 2> gpurun_out/r1c_bench_f64.err; echo "rc $?" >> gpurun_out/r1c_bench_f64.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1c_bench_ref.json 2> gpurun_out/r1c_bench_ref.err
python bench.py --precision f32 --no-cpu-baseline --steps 5 > gpurun_out/r1c_bench_f32.json 2> gpurun_out/r1c_bench_f32.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r1c_launches_f64.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1 --gl-steps 1 > gpurun_out/r1c_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:frontend_kernel --launch-skip 3 -c 1 -f -o gpurun_out/r1c_fe_f64 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --e2e-steps 1 --gl-batch 0 > gpurun_out/r1c_ncu_fe.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gl_kernel --launch-skip 150 -c 1 -f -o gpurun_out/r1c_gl python bench.py --steps 1 --warmup 3 --utts 500 --no-cpu-baseline --e2e-steps 1 > gpurun_out/r1c_ncu_gl.log 2>&1
ls -la gpurun_out
