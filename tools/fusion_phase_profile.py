"""Phase time line of fusion_node's CTA 0 (globaltimer stamps, ns): python tools/fusion_phase_profile.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from camouflage_multimodal_b200 import _lib, synth  # noqa: E402
from camouflage_multimodal_b200.pipeline import EvalPipeline  # noqa: E402

dev = torch.device("cuda")
gnn, det = bench.make_models()
kg, _ = bench.load_kg()
pipe = EvalPipeline(gnn.to(dev), det.to(dev), kg.to(dev), label_cap=bench.LABEL_CAP)
b = synth.make_batch(0, 32, seed=1234, device=dev).as_decoded()
for _ in range(3):
    pipe.run(b.images, b.labels, b.canny, b.gt)
torch.cuda.synchronize()
buf = torch.zeros(40 + 2 * 1024, dtype=torch.int64, device=dev)
lib = _lib.load()
lib.cm_debug_fusion_phase_stamps(_lib.ptr(buf))
pipe.run(b.images, b.labels, b.canny, b.gt)
torch.cuda.synchronize()
lib.cm_debug_fusion_phase_stamps(None)
_lib.timing_enable(True)
pipe.run(b.images, b.labels, b.canny, b.gt)
torch.cuda.synchronize()
print("event-timed (eager):", {k: round(v[1] * 1e3, 1) for k, v in _lib.timing_report().items() if k.startswith("fusion") or k.startswith("kg")}, "us")
_lib.timing_enable(False)
t = buf.cpu().tolist()
t0 = t[32]
names = ["r", "scores", "V2", "P1.VO", "hh0", "Y=hh0.W2a", "hh1", "Y+=hh1.W2b"]
print("phase          mma_issue  mma_go   epi_start  epi_done   (us after kernel start; MMA time ~= epi_start - mma_go)")
for ph in range(8):
    f = lambda v: f"{(v - t0) / 1e3:8.2f}" if v else "       -"
    print(f"{ph} {names[ph]:12s} {f(t[ph])} {f(t[8 + ph])} {f(t[16 + ph])} {f(t[24 + ph])}")
print("epi1 chunk loops done: half 0 (softmax -> P1) at", round((t[33] - t0) / 1e3, 2), " half 1 (S2 -> global) at", round((t[34] - t0) / 1e3, 2))
print("MODE 2 epi7: y tile written (before the column sums) at", round((t[39] - t0) / 1e3, 2))
print("MODE 2 kg pass (row 2 = its start / end): V2 half 0 staged / products done / V2 half 1 staged at", *(round((t[k] - t0) / 1e3, 2) for k in (35, 36, 37)))
print("epi1 first chunk (half 0): after TMEM load + bias", round((t[35] - t0) / 1e3, 2), " after softmax", round((t[36] - t0) / 1e3, 2), " after P1 store", round((t[37] - t0) / 1e3, 2))

# every CTA's start / end: launch ramp and tail of the grid
import numpy as np  # noqa: E402
ct = np.array(t[40:], dtype=np.int64).reshape(-1, 2)
ct = ct[ct[:, 0] > 0]
st, en = (ct[:, 0] - ct[:, 0].min()) / 1e3, (ct[:, 1] - ct[:, 0].min()) / 1e3
print("CTA 0: all warps done (before tcgen05.dealloc) at", round((t[38] - t0) / 1e3, 2), " after dealloc at", round((ct[0, 1] - t0) / 1e3, 2))
print("CTA 0: first instruction", round((ct[0, 0] - t0) / 1e3, 2), "us before its post-prologue stamp")
print(f"{len(ct)} CTAs: start min/median/max {st.min():.2f}/{np.median(st):.2f}/{st.max():.2f} us, end min/median/max "
      f"{en.min():.2f}/{np.median(en):.2f}/{en.max():.2f} us, tile time min/median/max {(en - st).min():.2f}/{np.median(en - st):.2f}/{(en - st).max():.2f} us")
