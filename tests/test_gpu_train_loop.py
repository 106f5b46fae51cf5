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


def test_eval_after_fused_adamw_uses_the_updated_weights():
    """cm_adamw_step writes parameters through raw pointers, so Tensor._version never moves; the
    caches of folded / converted weights (detector: cm_fusion_fold blob, GNN: cm_gnn_fold blob) must still refresh.
    Train-mode steps with FusedAdamW, then eval() output against the oracle on the UPDATED state dict."""
    from camouflage_multimodal_b200 import synth
    from camouflage_multimodal_b200.fusion_model import build_multimodal_model
    from camouflage_multimodal_b200.region_graph import RegionGraphGNN
    from camouflage_multimodal_b200.train_loop import FusedAdamW
    from oracle import fusion as ofus
    from oracle import gnn as ognn
    torch.manual_seed(5)
    det = build_multimodal_model({}).cuda()
    kg = synth.synthetic_kg(13).cuda()
    gen = torch.Generator().manual_seed(3)
    rg = torch.relu(torch.randn((2, 150, 128), generator=gen)).cuda()
    det.eval()
    with torch.no_grad():
        before = det(rg, kg[None].expand(2, -1, -1))[0].clone()          # builds the folded-weight cache
    opt = FusedAdamW(det.parameters(), lr=5e-2, weight_decay=1e-4, max_norm=1.0)
    for _ in range(3):
        opt.zero_grad()
        for p, g in zip(opt.params, opt.grads):
            g.copy_(torch.randn(p.shape, generator=gen).cuda())
        opt.step()
    with torch.no_grad():
        after = det(rg, kg[None].expand(2, -1, -1))[0]
    fsd = {k: v.detach().cpu() for k, v in det.state_dict().items()}
    with torch.no_grad():
        ref = ofus.detector_forward(fsd, rg.cpu(), kg.cpu()[None].expand(2, -1, -1))[0]
    assert float((after - before).abs().max()) > 1e-2                     # the step really moved the output
    np.testing.assert_allclose(after.cpu().numpy(), ref.numpy(), rtol=0, atol=1e-2)
    # the same for the GNN's folded weights (no BatchNorm buffer is touched between the two eval passes)
    gnn = RegionGraphGNN(15, 128, 2).cuda().eval()
    hb = synth.make_batch(0, 1, seed=9, n_regions=[120])
    from camouflage_multimodal_b200.region_graph import create_region_graph
    data, _ = create_region_graph(hb.images[0].numpy().astype(np.float64), segments=hb.labels[0].numpy(), edges=hb.canny[0].numpy())
    with torch.no_grad():
        g_before = gnn(data)[0].clone()
    gopt = FusedAdamW(gnn.parameters(), lr=5e-2, weight_decay=0.0, max_norm=None)
    for p, g in zip(gopt.params, gopt.grads):
        g.copy_(torch.randn(p.shape, generator=gen).cuda())
    gopt.step()
    with torch.no_grad():
        g_after = gnn(data)[0]
    gsd = {k: v.detach().cpu() for k, v in gnn.state_dict().items()}
    with torch.no_grad():
        r = ognn.rg_gnn_forward(gsd, data.x.cpu(), data.edge_index.cpu(), data.edge_attr.cpu())
    assert float((g_after - g_before).abs().max()) > 1e-3
    np.testing.assert_allclose(g_after.cpu().numpy(), r["mask_out"].numpy(), rtol=0, atol=2e-3)
