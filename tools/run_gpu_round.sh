timeout 900 python bench.py > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err
timeout 900 python bench.py --impl reference > gpurun_out/r2_bench_reference_arm.json 2> gpurun_out/r2_bench_reference_arm.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-torch-gpu-baseline --video-frames 0 > gpurun_out/r2_ncu_bench.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'gemm_bf16|flash_attn|window_attn|mlp_fused' -o gpurun_out/r2_top python tools/ncu_top_kernels.py > gpurun_out/r2_ncu_top.log 2>&1
timeout 900 python bench.py --sweep > gpurun_out/r2_config5_sweep_n1.json 2> gpurun_out/r2_sweep.err
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_smoke.txt 2>&1
