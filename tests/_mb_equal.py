"""Dev tool: bit checksum of the image-encoder outputs at batch 32 for the micro-batch given by SAM2B_MICROBATCH."""
import os, sys, torch
sys.path.insert(0, "/root/repo")
from oracle import sam2_oracle as O
from sam2_opt_b200.engine import Engine
e = Engine(O.make_weights(0))
g = torch.Generator(device="cuda").manual_seed(3)
img = torch.randn(32, 3, 1024, 1024, device="cuda", generator=g)
outs = e.image_encode(img, True, False)
print(os.environ.get("SAM2B_MICROBATCH"), [int(o.contiguous().view(torch.int32).long().sum().item()) for o in outs],
      [float(o[31].abs().max()) for o in outs])
