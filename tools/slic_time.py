"""Device SLIC: per-stage time at batch 32 and agreement with the CPU oracle on a few images."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from camouflage_multimodal_b200 import _lib, synth  # noqa: E402
from camouflage_multimodal_b200.region_graph import slic_segments  # noqa: E402
from oracle import slic as oslic  # noqa: E402

b = synth.make_batch(0, 32, seed=1234, device=torch.device("cuda")).as_decoded()
for _ in range(2):
    lab = slic_segments(b.images)
torch.cuda.synchronize()
_lib.timing_enable(True)
for _ in range(3):
    lab = slic_segments(b.images)
torch.cuda.synchronize()
rep = _lib.timing_report()
_lib.timing_enable(False)
for k, (c, ms) in sorted(rep.items(), key=lambda kv: -kv[1][1]):
    print(k, round(ms / 3, 3), "ms per batch of 32")
lab = lab.cpu().numpy()
for i in range(4):
    ref = oslic.slic(b.images[i].cpu().numpy())
    print("image", i, "agree", float((lab[i] == ref).mean()), "regions", len(np.unique(lab[i])), len(np.unique(ref)))
print("regions per image:", sorted(len(np.unique(l)) for l in lab))
