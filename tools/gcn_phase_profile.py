"""Phase stamps (%globaltimer, ns) of one CTA of the fused GCN layer kernels.  Needs a library built with
-DCM_GC_PROFILE (make -C camouflage_multimodal_b200/csrc EXTRA=-DCM_GC_PROFILE); not part of the product build."""
import ctypes
import sys

import torch

sys.path.insert(0, ".")
import bench  # noqa: E402
from camouflage_multimodal_b200 import _lib, synth  # noqa: E402
from camouflage_multimodal_b200.pipeline import EvalPipeline  # noqa: E402

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 32
dev = torch.device("cuda", 0)
lib = _lib.load()
gnn, det = bench.make_models()
kg, _ = bench.load_kg()
gnn, det, kg = gnn.to(dev), det.to(dev), kg.to(dev)
pipe = EvalPipeline(gnn, det, kg, label_cap=bench.LABEL_CAP, lane=0)
b = synth.make_batch(0, batch, seed=1234, device=dev).as_decoded()
for _ in range(5):
    pipe.run(b.images, b.labels, b.canny, b.gt)
torch.cuda.synchronize()
buf = (ctypes.c_ulonglong * 16)()
lib.cm_debug_gcn_profile.restype = ctypes.c_int
assert lib.cm_debug_gcn_profile(buf) == 0
t = list(buf)
names = ["start", "setup+edges", "gather", "mma0", "epi0", "mma1", "epi1", "mma2", "epi2", "end"]
print("last launch (the chained tail kernel), ns since CTA start:")
for i, n in enumerate(names):
    if t[i]:
        print(f"  {n:12s} {t[i] - t[0]:8d}")
