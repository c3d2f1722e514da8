"""Golden vectors for row f3 (`distCUDA2`): outputs of the REFERENCE's own simple_knn.cu (compiled in place for
sm_100 by `make -C oracle refknn` -> oracle/_ref/libref_knn.so) run on a B200.  Run on the GPU box:
    python tests/golden/make_golden_knn.py gpurun_out/knn_ref.npz
then copy the file to tests/golden/knn_ref.npz.  Seeded inputs are stored next to the outputs."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as orc  # noqa: E402

g = np.random.default_rng(11)
clouds = {
    "gauss": g.normal(size=(2000, 3)).astype(np.float32),
    "duplicates": np.concatenate([g.uniform(-1, 1, (1200, 3)).astype(np.float32)] * 1 + [np.zeros((3, 3), np.float32)]),
    "plane": np.concatenate([g.uniform(0, 1, (1500, 2)), np.zeros((1500, 1))], 1).astype(np.float32),
    "tiny3": np.array([[0, 0, 0], [1, 0, 0], [0, 2, 0]], np.float32),
}
clouds["duplicates"] = np.concatenate([clouds["duplicates"], clouds["duplicates"][:300]])
out = {}
for k, p in clouds.items():
    pts = torch.from_numpy(p).cuda()
    ref = torch.zeros(p.shape[0], device="cuda")
    orc.ref_knn3_mean_dist2(pts.data_ptr(), p.shape[0], ref.data_ptr())
    out["points_" + k] = p
    out["ref_" + k] = ref.cpu().numpy()
np.savez_compressed(sys.argv[1], **out)
print("wrote", sys.argv[1], {k: v.shape for k, v in out.items()})
