SAM2B_FUSED_MLP=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:mlp_fused --launch-skip 8 --launch-count 3 -o gpurun_out/r2s_mlp_fused -f python -c "
import os,sys,torch
sys.path.insert(0,'.')
from sam2_opt_b200 import engine as E
from sam2_opt_b200.synth import synth_image_norm
from sam2_opt_b200.synth_weights import make_weights
eng=E.Engine(make_weights(0),torch.device('cuda:0'))
img=synth_image_norm(16,seed=1).cuda()
for _ in range(2): eng.image_encode(img,True,False)
torch.cuda.synchronize()
" > gpurun_out/r2s_ncu.log 2>&1
tail -1 gpurun_out/r2s_ncu.log
