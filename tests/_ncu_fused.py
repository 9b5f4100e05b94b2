"""Dev tool: one image-encoder pass at batch 4 with the fused MLP on (for ncu -k regex:mlp_fused)."""
import os, sys, torch
os.environ["SAM2B_FUSED_MLP"] = "1"
sys.path.insert(0, "/root/repo")
from oracle import sam2_oracle as O
from sam2_opt_b200.engine import Engine
e = Engine(O.make_weights(0))
img = torch.randn(4, 3, 1024, 1024, device="cuda")
e.image_encode(img, True, False)
torch.cuda.synchronize()
print("ok")
