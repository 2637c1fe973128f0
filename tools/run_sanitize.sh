# usage: bash tools/run_sanitize.sh TAG -- compute-sanitizer memcheck over a subset of the GPU parity tests
TAG=${1:-x}
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 99 --log-file gpurun_out/${TAG}_memcheck.log python -m pytest tests -m gpu -x -q -k "tile_boundaries or prefetch_alignments or augment or specaugment or noise_fn or momentum or run_seams or long_form or loader_collate" > gpurun_out/${TAG}_memcheck_pytest.log 2>&1
echo "rc $?"; tail -3 gpurun_out/${TAG}_memcheck_pytest.log; tail -5 gpurun_out/${TAG}_memcheck.log
