set -x
nvidia-smi --query-gpu=name,memory.total --format=csv
free -g | head -2; nproc
( time python -m pytest tests -m gpu -x -q -s -p no:cacheprovider 2>&1 | tail -150 ) > gpurun_out/r2a_gpu_tests.log 2>&1
echo "pytest exit $?" >> gpurun_out/r2a_gpu_tests.log
( time python bench.py --steps 5 --warmup 3 ) > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err
( time python bench.py --impl reference --steps 2 --warmup 1 ) > gpurun_out/r2a_bench_ref.json 2> gpurun_out/r2a_bench_ref.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.log 2>&1
tail -3 gpurun_out/r2a_gpu_tests.log; tail -c 600 gpurun_out/r2a_bench.err; tail -2 gpurun_out/r2a_smoke.log
