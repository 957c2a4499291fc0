#!/usr/bin/env python
"""One fused-SSIM forward + backward at 1080p x 3 channels (the ncu target for csrc/ssim.cu)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import fused_ssim as fs  # noqa: E402

g = torch.Generator().manual_seed(0)
a = torch.rand(1, 3, 1080, 1920, generator=g).cuda().requires_grad_(True)
b = torch.rand(1, 3, 1080, 1920, generator=g).cuda()
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    a.grad = None
    fs.fused_ssim(a, b).backward()
torch.cuda.synchronize()
print("ok")
