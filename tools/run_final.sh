# usage: bash tools/run_final.sh TAG -- everything the round's evidence needs: GPU tests, both bench arms, launch list,
# full ncu captures of the two hot kernels
TAG=${1:-x}
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/${TAG}_pytest.log; tail -2 gpurun_out/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc $?"
python bench.py > gpurun_out/${TAG}_bench_f64.json 2> gpurun_out/${TAG}_bench_f64.err; echo "bench rc $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "ref rc $?"
python bench.py --precision f32 --no-cpu-baseline --steps 5 > gpurun_out/${TAG}_bench_f32.json 2> gpurun_out/${TAG}_bench_f32.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${TAG}_launches_f64.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1 --gl-steps 1 > gpurun_out/${TAG}_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:frontend_ --launch-skip 3 -c 1 -f -o gpurun_out/${TAG}_fe_f64 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --e2e-steps 1 --gl-batch 0 > gpurun_out/${TAG}_ncu_fe.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gl_ --launch-skip 150 -c 1 -f -o gpurun_out/${TAG}_gl python bench.py --steps 1 --warmup 3 --utts 500 --no-cpu-baseline --e2e-steps 1 > gpurun_out/${TAG}_ncu_gl.log 2>&1
echo done
