#!/usr/bin/env python
"""The multi-GPU fused all-reduce + Adam kernel with 8 emulated ranks on ONE GPU (cooperative
launch; the "peers" are local HBM, so this shows the instruction path — registers, occupancy,
stall mix — not link behaviour). The script ncu is pointed at:

    python bench/profile_fused_emulated.py && \
    ncu --set full --clock-control none --import-source on -k regex:fused_step -c 3 -o out python bench/profile_fused_emulated.py
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from fluxmpi_b200 import rules as R                                # noqa: E402
from fluxmpi_b200.ops import fused as F                            # noqa: E402
from fluxmpi_b200.parallel.device_world import EmulatedWorld       # noqa: E402

dev = torch.device("cuda", 0)
W, n = 8, 64 * 4096 * 8                      # 8 x 8 MB of fp32 gradients
ew = EmulatedWorld(W, dev, window_mb=80, stage_mb=1)
p_off, g_off = ew.heap_alloc(n * 4), ew.heap_alloc(n * 4)
st_off = ew.heap_alloc(2 * W * 64 * 4096 * 4)
for r in range(W):
    ew.window(r, p_off, n * 4).view(torch.float32).normal_()
    ew.window(r, g_off, n * 4).view(torch.float32).normal_()
h = F.hyper_for(R.Adam(1e-3), zero_grad=True)
m = [torch.zeros(n, device=dev) for _ in range(W)]
v = [torch.zeros(n, device=dev) for _ in range(W)]
step = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(W)]
for _ in range(3):
    ew.fused_step(g_off, p_off, torch.float32, m, v, None, 0, n, h, step)                       # slice-owner flavour
ew.fused_step_oneshot(g_off, p_off, torch.float32, m, v, None, 0, 64 * 4096, h, step, st_off, 2 * W * 64 * 4096 * 4)
torch.cuda.synchronize()
ew.check_health()
print("ok")
