timeout 900 python -m pytest tests/test_memory_attention_gpu.py tests/test_gemm_gpu.py tests/test_predictors_gpu.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -15 > gpurun_out/r3g_tests.log
timeout 300 python tools/video_profile.py 16 40 > gpurun_out/r3g_video_profile_16obj.txt 2>&1
timeout 300 python tools/video_profile.py 1 48 > gpurun_out/r3g_video_profile_1obj.txt 2>&1
