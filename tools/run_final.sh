# usage: bash tools/run_final.sh TAG -- everything a round's evidence needs: GPU tests, smoke, both bench arms, launch list,
# full ncu captures of the two hot kernels (each ncu run only after the same command has exited 0 without ncu)
TAG=${1:-x}
python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/${TAG}_pytest.log; tail -2 gpurun_out/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc $?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "ref rc $?"
python bench.py > gpurun_out/${TAG}_bench_f64.json 2> gpurun_out/${TAG}_bench_f64.err; echo "bench rc $?"
python bench.py --precision f32 --no-cpu-baseline --steps 5 --gl-batch 0 --longform-seconds 0 --loader-batch 0 --io-utts 0 > gpurun_out/${TAG}_bench_f32.json 2> gpurun_out/${TAG}_bench_f32.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1 --gl-steps 1 --io-utts 0"
$CMD > gpurun_out/${TAG}_plain_l.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_f64.csv $CMD > gpurun_out/${TAG}_ncu_l.log 2>&1
bash tools/run_prof.sh ${TAG} both
echo done
