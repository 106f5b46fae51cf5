"""Golden vectors for the oracle's AggressiveFocalLoss from the REFERENCE's own class (run in the build container only:
it reads /root/reference/models/multimodal/train_multimodal.py, extracts the class source with ast -- the module itself
imports tqdm / dataset code -- and executes it on seeded inputs).  Writes tests/golden/focal_loss.npz."""
import ast
import os

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

SRC = "/root/reference/models/multimodal/train_multimodal.py"
tree = ast.parse(open(SRC).read())
cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "AggressiveFocalLoss")
ns = {"torch": torch, "nn": nn, "F": F}
exec(compile(ast.Module(body=[cls], type_ignores=[]), SRC, "exec"), ns)
gen = torch.Generator().manual_seed(20240607)
x = (torch.randn((24, 2), generator=gen, dtype=torch.float64) * 2.5).requires_grad_(True)
t = (torch.rand(24, generator=gen) > 0.6).long()
loss = ns["AggressiveFocalLoss"](alpha=0.75, gamma=3.0)(x, t)
loss.backward()
out = os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "focal_loss.npz")
np.savez(out, logits=x.detach().numpy(), targets=t.numpy(), loss=loss.detach().numpy(), grad=x.grad.numpy())
print("wrote", out, float(loss))
