# usage: bash tools/run_sanitize.sh TOOL   (memcheck | racecheck | synccheck) -- ONE tool per gpurun call (B200_PROFILING.md)
# Runs the GPU parity subset that exercises seams, prefetch alignments, the momentum ring, linked seams, the masking and
# the tcgen05 prenet kernels under compute-sanitizer; the log tail goes to gpurun_out/sanitize_<tool>.log.
TOOL=${1:-memcheck}
K="rfft_core or irfft_core or ragged or tile or alignment or int16 or short or momentum or classic_zero or trim_and_minimum or sharded or augment or noise_fn or prenet or loader"
timeout 1500 compute-sanitizer --tool $TOOL --error-exitcode 99 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "$K" > gpurun_out/sanitize_${TOOL}.log 2>&1
echo "rc $?" >> gpurun_out/sanitize_${TOOL}.log
grep -E "ERROR SUMMARY|passed|failed|rc " gpurun_out/sanitize_${TOOL}.log | tail -5
