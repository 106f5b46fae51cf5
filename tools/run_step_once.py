"""One eager batch-32 step of the evaluation pipeline (for ncu captures): python tools/run_step_once.py [n_steps]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from camouflage_multimodal_b200 import synth  # noqa: E402
from camouflage_multimodal_b200.pipeline import EvalPipeline  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1
dev = torch.device("cuda")
gnn, det = bench.make_models()
kg, _ = bench.load_kg()
pipe = EvalPipeline(gnn.to(dev), det.to(dev), kg.to(dev), label_cap=bench.LABEL_CAP)
b = synth.make_batch(0, 32, seed=1234, device=dev).as_decoded()
for _ in range(2):
    pipe.run(b.images, b.labels, b.canny, b.gt)
torch.cuda.synchronize()
torch.cuda.profiler.start()
for _ in range(n):
    pipe.run(b.images, b.labels, b.canny, b.gt)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
