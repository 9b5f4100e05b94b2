"""Dev tool: image-encoder throughput at batch 32 for the micro-batch given by SAM2B_MICROBATCH."""
import os, sys, torch
sys.path.insert(0, "/root/repo")
from oracle import sam2_oracle as O
from sam2_opt_b200.engine import Engine
e = Engine(O.make_weights(0))
n = 32
img = torch.randn(n, 3, 1024, 1024, device="cuda")
for _ in range(2): e.image_encode(img, True, False)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(3): e.image_encode(img, True, False)
b.record(); torch.cuda.synchronize()
ms = a.elapsed_time(b) / 3
print(f"microbatch={os.environ.get('SAM2B_MICROBATCH')} n={n}: {ms:.2f} ms/step  {n/ms*1e3:.1f} img/s")
