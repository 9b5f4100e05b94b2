set -x
python -m pytest tests/test_memory_attention_gpu.py -m gpu -q -x -s -p no:cacheprovider 2>&1 | tail -30 > gpurun_out/r2f_memattn_tests.log
cat gpurun_out/r2f_memattn_tests.log
python -m pytest tests/test_predictors_gpu.py tests/test_video_long_gpu.py tests/test_speedup_reference_gpu.py tests/test_threads_gpu.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -15 > gpurun_out/r2f_pred_tests.log
cat gpurun_out/r2f_pred_tests.log
python tools/video_profile.py 1 40 2>&1 | head -14 > gpurun_out/r2f_video_prof_1.log
python tools/video_profile.py 16 40 2>&1 | head -14 > gpurun_out/r2f_video_prof_16.log
cat gpurun_out/r2f_video_prof_1.log gpurun_out/r2f_video_prof_16.log
