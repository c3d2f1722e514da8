"""Driver for the ncu capture of the "next row" kernels (f1-f4): one call of each at 1 M elements.
    ncu --set full --clock-control none -k regex:'k_adam_step|k_compact|k_knn3|k_knn_leaves|k_select_box|k_edit_apply|k_prune_near' \
        -o gpurun_out/next_rows python profiles/next_rows_driver.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import egr_b200  # noqa: E402
from egr_b200 import edit, optim, prune, scenes  # noqa: E402

n = 1_000_000
pts = torch.from_numpy(scenes.kitchen_scene(n, seed=0)["mean"]).cuda()
for _ in range(2):
    egr_b200.distCUDA2(pts)
names = optim.NAMES
params = {k: torch.randn(n, optim.CHANNELS[k], device="cuda") * 0.3 for k in names}
fused = optim.FusedAdam(params)
rec = torch.randn(n, 24, device="cuda")
for _ in range(2):
    fused.step(rec, {k: 1e-3 for k in names}, scale_decay=0.999, zero_grads=False)
keep = torch.rand(n, device="cuda") < 0.9
tensors = list(params.values()) + list(fused.exp_avg.values()) + list(fused.exp_avg_sq.values()) + [torch.randn(n, device="cuda") for _ in range(6)]
for _ in range(2):
    prune.compact(tensors, keep)
m = edit.select_box(pts, (0.5, 0.2, -1e30), (1.5, 1.4, 1.0))
edit.apply_edit(m, mean=pts.clone(), translate=(0.1, 0, 0), center=(1.0, 0.8, 0.5), scale_factor=1.1, rotate_deg=(0, 0, 30))
k8 = torch.ones(n, dtype=torch.uint8, device="cuda")
prune.drop_near_cameras(k8, pts, torch.rand(300, 3, device="cuda") * 4 - 2, torch.rand(300, device="cuda") * 0.1)
torch.cuda.synchronize()
print("ok")
