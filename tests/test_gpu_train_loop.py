"""region targets, the two training criteria and clip + AdamW against the CPU oracle / torch itself."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _graph(batch=3, seed=5):
    from camouflage_multimodal_b200 import synth
    from camouflage_multimodal_b200.region_graph import build_region_graphs
    b = synth.make_batch(0, batch, seed=seed, device=torch.device("cuda")).as_decoded()
    labels = b.labels.to(torch.int32)
    g = build_region_graphs(b.images, labels, b.canny, label_cap=640)
    return b, labels, g


def test_region_targets_match_reference_thresholds():
    from camouflage_multimodal_b200.train_loop import region_targets
    from oracle import train_loop as otl
    b, labels, g = _graph()
    gen = torch.Generator().manual_seed(3)
    # blobs of 0 / 255 with soft borders, so region means land on both sides of (and near) the thresholds
    masks = []
    for k in range(3):
        m = (torch.rand((labels.shape[0], 16, 16), generator=gen) > 0.55).float()
        m = torch.nn.functional.interpolate(m[:, None], size=labels.shape[1:], mode="bilinear")[:, 0]
        masks.append((m * 255).round().to(torch.uint8).cuda())
    y, iy, ey = region_targets(g, labels, *masks)
    ptr = g.node_ptr.cpu().tolist()
    for i in range(labels.shape[0]):
        oy, oiy, oey = otl.region_targets(labels[i].cpu().numpy(), *[m[i].cpu().numpy() for m in masks])
        assert y[ptr[i]:ptr[i + 1]].cpu().tolist() == oy
        assert iy[ptr[i]:ptr[i + 1]].cpu().tolist() == oiy
        assert ey[ptr[i]:ptr[i + 1]].cpu().tolist() == [float(v) for v in oey]
    assert 0 < sum(oy) < len(oy)


def test_rg_node_loss_value_and_gradients():
    from camouflage_multimodal_b200.train_loop import RegionGraphLoss
    from oracle import train_loop as otl
    gen = torch.Generator().manual_seed(11)
    n, cap = 777, 800
    mask = (torch.randn((cap, 2), generator=gen) * 2).cuda().requires_grad_(True)
    inst = (torch.randn((cap, 2), generator=gen) * 2).cuda().requires_grad_(True)
    edge = (torch.randn((cap, 1), generator=gen) * 3).cuda().requires_grad_(True)
    y = (torch.rand(cap, generator=gen) > 0.8).long().cuda()
    iy = (torch.rand(cap, generator=gen) > 0.6).long().cuda()
    ey = (torch.rand(cap, generator=gen) > 0.7).float().cuda()
    n_dev = torch.tensor([n], dtype=torch.int32, device="cuda")
    crit = RegionGraphLoss()
    loss = crit(mask, inst, edge, y, iy, ey, n_dev)
    loss.backward()
    om, oi, oe = [t.detach().cpu().double()[:n].requires_grad_(True) for t in (mask, inst, edge)]
    ol, *_ = otl.rg_node_loss(om, oi, oe, y.cpu()[:n], iy.cpu()[:n], ey.cpu()[:n])
    ol.backward()
    assert abs(float(loss.detach()) - float(ol.detach())) <= 2e-6 * abs(float(ol.detach()))
    for got, ref in ((mask.grad, om.grad), (inst.grad, oi.grad), (edge.grad, oe.grad)):
        np.testing.assert_allclose(got[:n].cpu().numpy(), ref.numpy(), rtol=2e-5, atol=1e-9)
        assert float(got[n:].abs().max()) == 0.0


def test_multimodal_loss_and_focal_loss():
    from camouflage_multimodal_b200.train_loop import AggressiveFocalLoss, MultimodalLoss
    from oracle import train_loop as otl
    gen = torch.Generator().manual_seed(17)
    G = 16
    mask = (torch.randn((G, 2), generator=gen) * 2).cuda().requires_grad_(True)
    inst = (torch.randn((G, 2), generator=gen) * 2).cuda().requires_grad_(True)
    edge = (torch.randn((G, 1), generator=gen) * 2).cuda().requires_grad_(True)
    score = torch.rand((G, 1), generator=gen).cuda().requires_grad_(True)
    ml = (torch.rand(G, generator=gen) > 0.5).long().cuda()
    el = (torch.rand(G, generator=gen) > 0.5).float().cuda()
    sl = torch.rand(G, generator=gen).cuda()
    loss = MultimodalLoss()(mask, inst, edge, score, ml, el, sl)
    loss.backward()
    o = [t.detach().cpu().double().requires_grad_(True) for t in (mask, inst, edge, score)]
    ol = otl.multimodal_loss(*o, ml.cpu(), el.cpu(), sl.cpu())
    ol.backward()
    assert abs(float(loss.detach()) - float(ol.detach())) <= 3e-6 * abs(float(ol.detach()))
    for got, ref in zip((mask, inst, edge, score), o):
        np.testing.assert_allclose(got.grad.cpu().numpy(), ref.grad.numpy(), rtol=5e-5, atol=2e-7)
    # the focal term alone, batch mean as the reference module reduces it
    z = (torch.randn((9, 2), generator=gen) * 3).cuda().requires_grad_(True)
    t = (torch.rand(9, generator=gen) > 0.4).long().cuda()
    fl = AggressiveFocalLoss()(z, t)
    fl.backward()
    zo = z.detach().cpu().double().requires_grad_(True)
    fo = otl.aggressive_focal_loss(zo, t.cpu())
    fo.backward()
    assert abs(float(fl.detach()) - float(fo.detach())) <= 3e-6 * abs(float(fo.detach())) + 1e-9
    np.testing.assert_allclose(z.grad.cpu().numpy(), zo.grad.numpy(), rtol=5e-5, atol=2e-7)


@pytest.mark.parametrize("max_norm", [1.0, None])
def test_fused_adamw_matches_torch_clip_and_adamw(max_norm):
    from camouflage_multimodal_b200.train_loop import FusedAdamW
    gen = torch.Generator().manual_seed(23)
    shapes = [(128, 60), (128,), (3, 64, 128), (1,), (257, 33)]
    ref_p = [torch.nn.Parameter(torch.randn(s, generator=gen)) for s in shapes]
    dev_p = [torch.nn.Parameter(p.detach().clone().cuda()) for p in ref_p]
    ref_opt = torch.optim.AdamW(ref_p, lr=1e-3, weight_decay=1e-4)
    opt = FusedAdamW(dev_p, lr=1e-3, weight_decay=1e-4, max_norm=max_norm)
    for step in range(4):
        grads = [torch.randn(s, generator=gen) * (3.0 if step % 2 == 0 else 0.01) for s in shapes]
        for p, g in zip(ref_p, grads):
            p.grad = g.clone()
        for p, g in zip(dev_p, grads):
            p.grad.copy_(g.cuda())
        norm_ref = torch.nn.utils.clip_grad_norm_(ref_p, max_norm) if max_norm is not None else None
        ref_opt.step()
        opt.step()
        if norm_ref is not None:
            assert abs(float(opt.total_norm) - float(norm_ref)) <= 2e-6 * float(norm_ref)
            for p, q in zip(dev_p, ref_p):   # clip_grad_norm_ scales .grad in place
                np.testing.assert_allclose(p.grad.cpu().numpy(), q.grad.numpy(), rtol=3e-6, atol=1e-9)
        for p, q in zip(dev_p, ref_p):
            np.testing.assert_allclose(p.detach().cpu().numpy(), q.detach().numpy(), rtol=2e-6, atol=2e-7)
