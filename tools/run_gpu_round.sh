timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_image_encoder_gpu.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -3
python tests/_attn_bench.py 2>&1 | grep -E "ws=16|ws=64" | grep -E "B=16" | tee gpurun_out/r3a_attn_bench.log
