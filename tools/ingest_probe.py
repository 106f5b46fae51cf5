"""Host -> device staging rates on this box: copy engine (cudaMemcpyAsync from pinned memory) against cm_ingest_copy (SM loads
from the same pinned memory), idle and while device-resident steps keep the GPU busy.  python tools/ingest_probe.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from camouflage_multimodal_b200 import _lib, synth  # noqa: E402
from camouflage_multimodal_b200.pipeline import EvalPipeline  # noqa: E402

dev = torch.device("cuda")
lib = _lib.load()
n = 11010048
hosts = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(8)]
for h in hosts:
    h.random_(0, 255)
d = torch.empty(n, dtype=torch.uint8, device=dev)
gnn, det = bench.make_models()
kg, _ = bench.load_kg()
pipes = [EvalPipeline(gnn.to(dev), det.to(dev), kg.to(dev), label_cap=bench.LABEL_CAP, lane=i) for i in range(4)]
b = synth.make_batch(0, 32, seed=1234, device=dev).as_decoded()
graphs = [p.capture(b.images, b.labels, b.canny, b.gt)[0] for p in pipes]
lanes = [torch.cuda.Stream(dev) for _ in range(4)]
cs = torch.cuda.Stream(dev)


def busy(k):
    for j in range(k):
        with torch.cuda.stream(lanes[j % 4]):
            graphs[j % 4].replay()


def rate(fn, load):
    torch.cuda.synchronize()
    if load:
        busy(64)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(cs):
        e0.record(cs)
        for j in range(20):
            fn(hosts[j % 8])
        e1.record(cs)
    if load:
        busy(160)
    torch.cuda.synchronize()
    return 20 * n / (e0.elapsed_time(e1) * 1e-3) / 1e9


def ce(h):
    d.copy_(h, non_blocking=True)


def sm(ctas):
    def f(h):
        _lib.check(lib.cm_ingest_copy(h.data_ptr(), _lib.ptr(d), n, ctas, _lib.stream_ptr(dev)), "ingest")
    return f


ce(hosts[0]); sm(32)(hosts[0]); torch.cuda.synchronize()
assert torch.equal(d.cpu(), hosts[0])
for name, fn in [("copy engine", ce)] + [(f"SM copy, {c} CTAs", sm(c)) for c in (8, 16, 32, 64, 148)]:
    print(f"{name:22s} idle {rate(fn, False):6.1f} GB/s   under load {rate(fn, True):6.1f} GB/s")
