timeout 300 python tests/_attn_bench.py > gpurun_out/r3d_attn_bench.log 2>&1
