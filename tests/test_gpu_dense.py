"""GPU parity of the tcgen05/TMEM/TMA dense layer (cm_linear_f16) against a plain PyTorch fp32 reference of the
same op on the same fp16-rounded operands.  fp32 accumulation => tolerance is fp32 summation noise (1e-3 abs at
K=512 with O(1) outputs is generous); fp16 outputs additionally carry one fp16 rounding (2^-11 relative)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(m, n, k, *, bias=True, alpha=1.0, relu=False, residual=False, ln=False, live=None, lda=None, seed=0):
    from camouflage_multimodal_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(seed)
    lda = lda or k
    a = (torch.randn(m, lda, generator=g) * 0.5).to(torch.float16).cuda()
    w = (torch.randn(n, k, generator=g) * (1.0 / k ** 0.5)).to(torch.float16).cuda()
    b = torch.randn(n, generator=g).cuda() if bias else None
    r = (torch.randn(m, n, generator=g)).to(torch.float16).cuda() if residual else None
    gam = (1 + 0.1 * torch.randn(n, generator=g)).cuda() if ln else None
    bet = (0.1 * torch.randn(n, generator=g)).cuda() if ln else None
    out16 = torch.full((m, n), 7.0, dtype=torch.float16, device="cuda")
    out32 = torch.full((m, n), 7.0, dtype=torch.float32, device="cuda")
    mdev = torch.tensor([live], dtype=torch.int32, device="cuda") if live is not None else None
    p = _lib.ptr
    _lib.check(lib.cm_linear_f16(p(a), lda, p(w), m, n, k, p(mdev), p(b), float(alpha), int(relu), p(r), n, p(gam), p(bet),
                                  p(out16), p(out32), _lib.stream_ptr()), "cm_linear_f16")
    torch.cuda.synchronize()
    ref = a[:, :k].float() @ w.float().t()
    if bias:
        ref = ref + b
    ref = ref * alpha
    if residual:
        ref = ref + r.float()
    if relu:
        ref = ref.relu()
    if ln:
        ref = torch.nn.functional.layer_norm(ref, (n,), gam, bet, 1e-5)
    rows = live if live is not None else m
    err32 = (out32[:rows] - ref[:rows]).abs().max().item()
    err16 = (out16[:rows].float() - ref[:rows]).abs().max().item()
    scale = ref[:rows].abs().max().item()
    assert err32 <= 2e-3, (m, n, k, err32)
    assert err16 <= 2e-3 + scale * 2 ** -11, (m, n, k, err16)
    if rows < m:   # rows past the live count are never written
        assert (out32[rows:] == 7.0).all() and (out16[rows:].float() == 7.0).all()


@pytest.mark.parametrize("m,n,k", [(128, 128, 64), (300, 256, 128), (1000, 512, 256), (481, 256, 512), (77, 192, 128),
                                   (13, 256, 256), (15466, 128, 128), (200, 128, 72), (32, 2 * 256, 512), (5, 64, 256)])
def test_shapes(m, n, k):
    _run(m, n, k)


def test_epilogues():
    _run(500, 256, 256, relu=True)
    _run(500, 256, 256, alpha=0.1767767, bias=True)
    _run(500, 256, 256, residual=True)
    _run(500, 256, 256, residual=True, ln=True)
    _run(333, 128, 128, residual=True, ln=True, relu=False)
    _run(400, 256, 128, live=257)
    _run(400, 256, 64, live=1, lda=64)
    _run(260, 128, 64, lda=72, bias=False)
    _run(640, 512, 256, relu=True, live=600)


@pytest.mark.parametrize("m,n,k,scale,ta", [(1000, 256, 256, 1.0, 0), (300, 128, 300, 1.0, 0), (256, 256, 4000, 1.0, 1), (256, 128, 7000, 1e-5, 1),
                                            (512, 256, 236, 1e-7, 1), (777, 192, 60, 1.0, 0), (37, 5, 19, 1.0, 0)])
def test_training_gemm_building_block(m, n, k, scale, ta):
    """cm_train_gemm_f32 (the dense block of the training passes) against float64: large shapes run on tcgen05 with 16-bit
    thirds (fp16 parts for the forward pattern, bf16 parts + split-K for the weight-gradient pattern, whose operand may be
    tiny), small ones on fp32 tiles; either way <= 5e-5 of the output scale."""
    from camouflage_multimodal_b200 import _lib
    lib = _lib.load()
    gen = torch.Generator().manual_seed(m + k)
    a = (torch.randn((m, k), generator=gen) * scale).cuda()
    w = torch.randn((n, k), generator=gen).cuda()
    b = (torch.randn(n, generator=gen) * scale).cuda()
    c = torch.full((m, n), float("nan"), device="cuda")
    aa = a.t().contiguous() if ta else a
    _lib.check(lib.cm_train_gemm_f32(_lib.ptr(aa), _lib.ptr(w), _lib.ptr(b), m, n, k, ta, _lib.ptr(c), _lib.stream_ptr(a.device)), "cm_train_gemm_f32")
    ref = a.double() @ w.double().t() + b.double()
    assert torch.isfinite(c).all()
    assert float((c.double() - ref).abs().max() / ref.abs().max()) <= 5e-5


_SPLIT_CASES = [(1000, 256, 256, 0), (300, 128, 300, 0), (256, 256, 4000, 1), (512, 256, 236, 1), (640, 200, 1036, 1)]


def _split_case_outputs():
    from camouflage_multimodal_b200 import _lib
    lib = _lib.load()
    outs = []
    for m, n, k, ta in _SPLIT_CASES:
        gen = torch.Generator().manual_seed(3 * m + k)
        a = torch.randn((m, k), generator=gen).cuda()
        w = torch.randn((n, k), generator=gen).cuda()
        c = torch.full((m, n), float("nan"), device="cuda")
        aa = a.t().contiguous() if ta else a
        _lib.check(lib.cm_train_gemm_f32(_lib.ptr(aa), _lib.ptr(w), None, m, n, k, ta, _lib.ptr(c), _lib.stream_ptr(a.device)), "cm_train_gemm_f32")
        outs.append(c.cpu())
    return outs


def test_training_gemm_vector_operand_split_equals_the_scalar_split(tmp_path):
    """The 8-wide operand-split kernels (train_common.cu) write the same 16-bit thirds as the one-element kernels: the GEMM
    results of a process run with CM_TRAIN_SPLIT_SCALAR=1 are bit-equal to this process's."""
    import os
    import subprocess
    import sys
    here = _split_case_outputs()
    path = tmp_path / "scalar.pt"
    code = ("import sys, torch; sys.path.insert(0, %r); sys.path.insert(0, %r); import test_gpu_dense as t; "
            "torch.save(t._split_case_outputs(), %r)") % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                          os.path.dirname(os.path.abspath(__file__)), str(path))
    subprocess.run([sys.executable, "-c", code], check=True, env=dict(os.environ, CM_TRAIN_SPLIT_SCALAR="1"), timeout=600)
    there = torch.load(path)
    for (m, n, k, ta), x, y in zip(_SPLIT_CASES, here, there):
        assert torch.isfinite(x).all()
        assert torch.equal(x, y), (m, n, k, ta)
