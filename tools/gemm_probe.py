"""Precision probe of the training GEMM building block (cm_train_gemm_f32) against float64: python tools/gemm_probe.py"""
import sys
import os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from camouflage_multimodal_b200 import _lib
lib = _lib.load()
torch.manual_seed(0)
for (m, n, k, scale, ta) in [(1000, 256, 256, 1.0, 0), (1000, 256, 256, 1e-6, 0), (256, 256, 4000, 1.0, 1), (256, 128, 7000, 1e-5, 1), (777, 192, 60, 1.0, 0),
                             (236, 128, 128, 1.0, 0), (236, 64, 128, 1.0, 0), (300, 128, 300, 1.0, 0), (128, 128, 300, 1e-4, 1), (64, 128, 600, 1.0, 1),
                             (512, 256, 236, 1.0, 1), (236, 512, 256, 1.0, 0), (236, 256, 512, 1.0, 0)]:
    a = torch.randn(m, k, device="cuda") * scale
    w = torch.randn(n, k, device="cuda")
    b = torch.randn(n, device="cuda") * scale
    c = torch.full((m, n), float("nan"), device="cuda")
    aa = a.t().contiguous() if ta else a
    _lib.check(lib.cm_train_gemm_f32(_lib.ptr(aa), _lib.ptr(w), _lib.ptr(b), m, n, k, ta, _lib.ptr(c), _lib.stream_ptr(a.device)), "gemm")
    ref = a.double() @ w.double().t() + b.double()
    err = (c.double() - ref).abs().max().item() / ref.abs().max().item()
    print(m, n, k, scale, ta, "rel err", err)
