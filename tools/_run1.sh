set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_sam_heads_gpu.py -x -q -s > gpurun_out/heads_test.log 2>&1; echo "rc=$?" >> gpurun_out/heads_test.log
tail -30 gpurun_out/heads_test.log
timeout 400 python tests/_video_fps.py 64 > gpurun_out/fps_native.log 2>&1; tail -4 gpurun_out/fps_native.log
VP_NATIVE=0 timeout 400 python tests/_video_fps.py 64 > gpurun_out/fps_torch.log 2>&1; tail -4 gpurun_out/fps_torch.log
timeout 300 python tools/video_profile.py 16 40 > gpurun_out/vprof16_native.log 2>&1; head -30 gpurun_out/vprof16_native.log
timeout 300 python tools/video_profile.py 1 48 > gpurun_out/vprof1_native.log 2>&1; head -40 gpurun_out/vprof1_native.log
