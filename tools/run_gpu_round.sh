for pp in 0 1; do
SAM2B_ATTN_PINGPONG=$pp SAM2B_ATTN_TRACE=1 timeout 300 python -c "
import sys,torch
sys.path.insert(0,'.')
from sam2_opt_b200 import ops
g=torch.Generator(device='cuda').manual_seed(0)
qkv=torch.randn(4,64,64,3*576,device='cuda',generator=g).bfloat16()
for _ in range(2): ops.window_attention(qkv,8,64,False)
torch.cuda.synchronize()
" 2> gpurun_out/r3j_attn_trace_global_pp$pp.log
done
