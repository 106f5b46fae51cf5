"""GPU parity of the large-KG attention path (BASELINE configs[4]): the tcgen05 fused QK^T.softmax.V kernel against a
float64 softmax on the same fp16-rounded operands, and the whole fusion with a 1 000-row shared knowledge graph
against the oracle.  Tolerances: context rows 2e-3 abs (probabilities travel as fp16 into the PV product);
mask logits 5e-3 (half the north star's 1e-2 for reduced-precision logits)."""
import math

import numpy as np
import pytest
import torch

from camouflage_multimodal_b200 import synth

pytestmark = pytest.mark.gpu


def _ref_attention(q, k, v, heads, scale, segs):
    """float64 reference: per segment [a, b) of keys, per head of 32 dims."""
    q, k, v = q.double(), k.double(), v.double()
    mq, d = q.shape
    hd = d // heads
    outs = []
    for a, b in segs:
        o = torch.zeros((mq, d), dtype=torch.float64)
        if b > a:
            for h in range(heads):
                sl = slice(h * hd, (h + 1) * hd)
                p = torch.softmax(q[:, sl] @ k[a:b, sl].t() * scale, dim=-1)
                o[:, sl] = p @ v[a:b, sl]
        outs.append(o)
    return torch.cat(outs)


def _run_flash(q, k, v, seg_ptr, scale):
    from camouflage_multimodal_b200 import _lib
    lib = _lib.load()
    mq, d = q.shape
    mk = k.shape[0]
    ldv = (mk + 7) // 8 * 8
    vt = torch.zeros((d, ldv), dtype=torch.float16, device="cuda")
    vt[:, :mk] = v.t()
    n_seg = seg_ptr.numel() - 1 if seg_ptr is not None else 1
    out = torch.full((n_seg * mq, d), float("nan"), dtype=torch.float16, device="cuda")
    _lib.check(lib.cm_flash_attention_f16(_lib.ptr(q), mq, _lib.ptr(k), mk, _lib.ptr(vt), ldv, d, _lib.ptr(seg_ptr), n_seg, scale,
                                          _lib.ptr(out), d, _lib.stream_ptr()), "cm_flash_attention_f16")
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("mq,mk", [(200, 300), (128, 128), (57, 1000), (513, 77)])
def test_flash_attention_single_segment(mq, mk):
    g = torch.Generator().manual_seed(mq * 1000 + mk)
    q = (torch.randn((mq, 256), generator=g) * 1.5).half().cuda()
    k = (torch.randn((mk, 256), generator=g) * 1.5).half().cuda()
    v = torch.randn((mk, 256), generator=g).half().cuda()
    scale = 1.0 / math.sqrt(32)
    out = _run_flash(q, k, v, None, scale)
    ref = _ref_attention(q.cpu(), k.cpu(), v.cpu(), 8, scale, [(0, mk)])
    assert torch.isfinite(out).all()
    np.testing.assert_allclose(out.float().cpu().numpy(), ref.numpy(), rtol=0, atol=2e-3)


def test_flash_attention_ragged_segments():
    """Key segments that start at unaligned rows, a one-key segment and an empty one (zeros by definition here)."""
    g = torch.Generator().manual_seed(5)
    mq, mk = 150, 700
    q = (torch.randn((mq, 256), generator=g) * 2).half().cuda()
    k = (torch.randn((mk, 256), generator=g) * 2).half().cuda()
    v = torch.randn((mk, 256), generator=g).half().cuda()
    bounds = [0, 130, 131, 131, 500, 700]
    seg = torch.tensor(bounds, dtype=torch.int32, device="cuda")
    scale = 1.0 / math.sqrt(32)
    out = _run_flash(q, k, v, seg, scale)
    ref = _ref_attention(q.cpu(), k.cpu(), v.cpu(), 8, scale, list(zip(bounds[:-1], bounds[1:])))
    assert torch.isfinite(out).all()
    np.testing.assert_allclose(out.float().cpu().numpy(), ref.numpy(), rtol=0, atol=2e-3)
    # one key: the context row IS that value row
    assert torch.equal(out[mq:2 * mq], v[130].expand(mq, -1))


def test_fusion_large_kg_matches_oracle():
    """Whole CrossAttentionFusion + heads with 1 000 shared KG rows on a ragged batch (flash path) vs the oracle."""
    from camouflage_multimodal_b200.fusion_model import build_multimodal_model
    from oracle import fusion as ofus
    torch.manual_seed(3)
    det = build_multimodal_model({}).cuda().eval()
    gen = torch.Generator().manual_seed(17)
    sizes = [300, 77, 481]
    M, G, K = sum(sizes), 3, 1000
    rows = (torch.randn((M, 128), generator=gen).clamp_min(0) * 0.5).cuda()
    kg = synth.synthetic_kg(K).cuda()
    ptr = torch.tensor(np.concatenate([[0], np.cumsum(sizes)]), dtype=torch.int32, device="cuda")
    with torch.no_grad():
        res = det.forward_ragged(rows, ptr, G, kg[None])
    sd = {k: v.detach().cpu() for k, v in det.state_dict().items()}
    p = ptr.cpu().tolist()
    for g_i in range(G):
        with torch.no_grad():
            mask, inst, edge, score, _, fused = ofus.detector_forward(sd, rows[p[g_i]:p[g_i + 1]].cpu()[None], kg.cpu()[None])
        assert (res["mask_out"][g_i].cpu() - mask[0]).abs().max() <= 5e-3
        assert (res["score_out"][g_i].cpu() - score[0]).abs().max() <= 5e-3
        assert (res["fused"][g_i].cpu() - fused[0]).abs().max() <= 1e-2
    # the same call with attention maps requested takes the fp32 path and must agree
    with torch.no_grad():
        res2 = det.forward_ragged(rows, ptr, G, kg[None], return_attention=True)
    assert (res2["mask_out"] - res["mask_out"]).abs().max() <= 5e-3
