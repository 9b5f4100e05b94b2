timeout 300 python -m pytest tests/test_image_encoder_gpu.py -m gpu -q -x -p no:cacheprovider -k fused -s 2>&1 | tail -3
timeout 300 python tools/fused_mlp_probe.py 32 2>&1 | tee gpurun_out/r2x_fused_mlp_probe.log
