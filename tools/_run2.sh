set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multi_device_gpu.py -x -q > gpurun_out/t_2gpu.log 2>&1; echo "rc=$?" >> gpurun_out/t_2gpu.log
tail -3 gpurun_out/t_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench rc=$?"
tail -c 1500 gpurun_out/bench_n2.json
