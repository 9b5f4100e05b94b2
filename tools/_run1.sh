set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/ -x -q -m gpu > gpurun_out/gpu_tests.log 2>&1; echo "rc=$?" >> gpurun_out/gpu_tests.log
tail -15 gpurun_out/gpu_tests.log
timeout 400 python tests/_video_fps.py 64 > gpurun_out/fps_native.log 2>&1; tail -4 gpurun_out/fps_native.log
timeout 300 python tools/video_profile.py 1 48 > gpurun_out/vprof1_native.log 2>&1; head -12 gpurun_out/vprof1_native.log
