timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_memory_attention_gpu.py tests/test_image_encoder_gpu.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -4
python tests/_attn_bench.py 2>&1 | grep -E "ws=16|ws=64" | grep -E "B=16" | tee gpurun_out/r2p_attn_bench.log
python tools/video_profile.py 16 40 2>&1 | head -5
python tools/fused_mlp_probe.py 32 2>&1 | grep -E "SAM2B"
