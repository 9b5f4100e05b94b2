set -x
( time python -m pytest tests -m gpu -q -s -p no:cacheprovider 2>&1 | tail -120 ) > gpurun_out/r2b_gpu_tests.log 2>&1
python tools/sanitize_run.py > gpurun_out/r2b_plain_run.log 2>&1
( time timeout 600 compute-sanitizer --tool memcheck --print-limit 20 python tools/sanitize_run.py ) > gpurun_out/r2b_memcheck.log 2>&1
( time timeout 600 compute-sanitizer --tool racecheck --print-limit 20 python tools/sanitize_run.py ) > gpurun_out/r2b_racecheck.log 2>&1
python tools/ncu_top_kernels.py > gpurun_out/r2b_ncu_plain.log 2>&1 && \
( time timeout 900 ncu --set full --clock-control none --import-source on -k regex:'gemm_bf16|flash_attn|window_attn' -o gpurun_out/r2b_top -f python tools/ncu_top_kernels.py ) > gpurun_out/r2b_ncu.log 2>&1
ls -la gpurun_out/*.ncu-rep
tail -5 gpurun_out/r2b_gpu_tests.log; tail -5 gpurun_out/r2b_memcheck.log; tail -5 gpurun_out/r2b_racecheck.log; tail -3 gpurun_out/r2b_ncu.log
