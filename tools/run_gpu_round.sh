python tools/heads_profile.py 16 2>&1 | tail -50 | tee gpurun_out/r2j_heads_profile_16.log
