timeout 600 python -m pytest tests/test_image_encoder_gpu.py tests/test_gemm_gpu.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -4
timeout 300 python tools/fused_mlp_probe.py 32 2>&1 | tee gpurun_out/r2o_fused_mlp_probe.log
