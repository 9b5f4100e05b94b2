set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_graph_track_gpu.py -x -q -s > gpurun_out/graph_tests.log 2>&1; echo "rc=$?" >> gpurun_out/graph_tests.log
tail -8 gpurun_out/graph_tests.log
timeout 400 python tests/_video_fps.py 64 > gpurun_out/fps_graph.log 2>&1; tail -4 gpurun_out/fps_graph.log
VP_GRAPH=0 timeout 400 python tests/_video_fps.py 64 > gpurun_out/fps_nograph.log 2>&1; tail -4 gpurun_out/fps_nograph.log
timeout 400 python tests/_video_fps.py 128 > gpurun_out/fps_graph128.log 2>&1; tail -4 gpurun_out/fps_graph128.log
VP_GRAPH=0 timeout 400 python tests/_video_fps.py 128 > gpurun_out/fps_nograph128.log 2>&1; tail -4 gpurun_out/fps_nograph128.log
