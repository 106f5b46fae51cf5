"""GPU parity of the training step (BASELINE configs[3]): RegionGraphGNN forward under model.train() and the
gradients of every parameter through cm_gnn_backward, against torch autograd over the float64 oracle on the same
batched graph, loss and dropout masks.  Tolerances: outputs 1e-4 abs; each gradient tensor 2e-4 of its own
max-abs (+1e-6) -- fp32 kernels vs a float64 reference."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from camouflage_multimodal_b200 import synth

pytestmark = pytest.mark.gpu


def _loss(mask_out, inst_out, edge_out, y, iy, ey):
    """The reference's weighted multi-task loss (models/region_graph/train.py:325-353, task weights 1/1/1)."""
    dt, dev = mask_out.dtype, mask_out.device
    lm = F.cross_entropy(mask_out, y, weight=torch.tensor([1.0, 5.0], dtype=dt, device=dev))
    li = F.cross_entropy(inst_out, iy, weight=torch.tensor([1.0, 4.0], dtype=dt, device=dev))
    le = F.binary_cross_entropy_with_logits(edge_out.squeeze(-1), ey.to(dt), pos_weight=torch.tensor([3.0], dtype=dt, device=dev))
    return lm + li + le


def _batch(n_img=3, first=60):
    from camouflage_multimodal_b200.region_graph import batch_to_data, build_region_graphs
    b = synth.make_batch(first, n_img, seed=1234, n_regions=[140, 97, 201][:n_img]).to("cuda")
    g = build_region_graphs(b.images, b.labels, b.canny, label_cap=256)
    data = batch_to_data(g)
    n = data.x.shape[0]
    gen = torch.Generator().manual_seed(11)
    y = (torch.rand(n, generator=gen) < 0.3).long()
    iy = (torch.rand(n, generator=gen) < 0.4).long()
    ey = (torch.rand(n, generator=gen) < 0.25).float()
    return data, y, iy, ey


def _model(seed=3):
    from camouflage_multimodal_b200.region_graph import RegionGraphGNN
    torch.manual_seed(seed)
    m = RegionGraphGNN(15, 128, 2)
    with torch.no_grad():   # non-trivial BN affine parameters and conv biases
        for bn in (m.bn1, m.bn2, m.bn3, m.bn4):
            bn.weight.uniform_(0.5, 1.5)
            bn.bias.normal_(0, 0.2)
        for c in (m.conv1, m.conv2, m.conv3, m.conv4):
            c.bias.normal_(0, 0.1)
    return m.cuda()


def _oracle_grads(m, data, y, iy, ey, masks):
    from oracle import gnn as ognn
    sd = {k: v.detach().cpu().double().requires_grad_(v.dtype.is_floating_point and "running" not in k)
          for k, v in m.state_dict().items() if "num_batches" not in k}
    drop = None if masks is None else [None if t is None else t.cpu() for t in masks]
    r = ognn.rg_gnn_forward(sd, data.x.cpu().double(), data.edge_index.cpu(), data.edge_attr.cpu().double(),
                            training_bn=True, drop=drop)
    loss = _loss(r["mask_out"], r["instance_out"], r["edge_out"], y, iy, ey)
    loss.backward()
    return r, loss, {k: v.grad for k, v in sd.items() if v.requires_grad}


def _check(m, data, y, iy, ey, masks, outs, loss):
    r, oloss, og = _oracle_grads(m, data, y, iy, ey, masks)
    for got, key in zip(outs, ("mask_out", "instance_out", "edge_out")):
        np.testing.assert_allclose(got.detach().cpu().numpy(), r[key].detach().numpy(), rtol=0, atol=1e-4)
    assert abs(loss.item() - oloss.item()) <= 1e-5
    worst = {}
    for name, prm in m.named_parameters():
        ref = og[name].numpy()
        got = prm.grad.detach().cpu().numpy().astype(np.float64)
        assert got.shape == ref.shape, name
        tol = 2e-4 * np.abs(ref).max() + 1e-6
        worst[name] = np.abs(got - ref).max() / tol
    bad = {k: v for k, v in worst.items() if v > 1.0}
    assert not bad, f"gradient mismatch (error / tolerance): {bad}"


def test_gnn_train_step_gradients_match_autograd_oracle():
    data, y, iy, ey = _batch()
    m = _model().train()
    m.p_conv = m.p_fc = 0.0
    rm0 = m.bn1.running_mean.clone()
    outs = m(data)
    loss = _loss(*outs, y.cuda(), iy.cuda(), ey.cuda())
    loss.backward()
    _check(m, data, y, iy, ey, None, outs, loss)
    # BatchNorm running statistics moved like nn.BatchNorm1d's (momentum 0.1, unbiased variance)
    assert not torch.equal(rm0, m.bn1.running_mean) and int(m.bn1.num_batches_tracked) == 1


def test_gnn_train_step_with_dropout_masks():
    """Dropout through explicit keep masks: same masks on both sides, p = 0.3 / 0.2 as in the reference."""
    from camouflage_multimodal_b200 import training
    data, y, iy, ey = _batch(2, first=75)
    m = _model(5).train()
    n = data.x.shape[0]
    gen = torch.Generator().manual_seed(21)
    masks = [(torch.rand((n, 128 if k < 4 else 64), generator=gen) >= (0.3 if k < 3 else 0.2)).to(torch.uint8).cuda() for k in range(7)]
    row_ptr, col, w, self_w = m._csr_of(data)
    outs = training.GnnTrainFunction.apply(m, data.x, row_ptr, col, w, self_w, int(row_ptr[-1]), masks, *training._gnn_param_list(m))
    loss = _loss(*outs[:3], y.cuda(), iy.cuda(), ey.cuda())
    loss.backward()
    _check(m, data, y, iy, ey, masks, outs[:3], loss)


def test_gnn_train_step_is_reproducible_and_optimizer_step_reduces_loss():
    """No float atomics in the backward: two identical steps give bit-identical gradients; a few AdamW steps of the
    reference's loop (clip_grad_norm_ 1.0, train.py:355-357) reduce the loss."""
    data, y, iy, ey = _batch(2, first=81)
    grads = []
    for _ in range(2):
        m = _model(7).train()
        m.p_conv = m.p_fc = 0.0
        loss = _loss(*m(data), y.cuda(), iy.cuda(), ey.cuda())
        loss.backward()
        grads.append([p.grad.clone() for p in m.parameters()])
    assert all(torch.equal(a, b) for a, b in zip(*grads))
    m = _model(7).train()
    opt = torch.optim.AdamW(m.parameters(), lr=1e-3, weight_decay=1e-4)
    losses = []
    for _ in range(8):
        opt.zero_grad()
        loss = _loss(*m(data), y.cuda(), iy.cuda(), ey.cuda())
        loss.backward()
        torch.nn.utils.clip_grad_norm_(m.parameters(), max_norm=1.0)
        opt.step()
        losses.append(loss.item())
    assert losses[-1] < losses[0]
    m.eval()
    with torch.no_grad():
        mask, _, _ = m(data)     # the inference kernels pick up the updated weights and running statistics
    assert torch.isfinite(mask).all()


# ----------------------------------------------------------------------------------------------- fusion
def _fusion_loss(mask, inst, edge, score, y, iy, ey, sy):
    dt = mask.dtype
    return (F.cross_entropy(mask, y) + F.cross_entropy(inst, iy) + F.binary_cross_entropy_with_logits(edge.squeeze(-1), ey.to(dt))
            + F.binary_cross_entropy(score.squeeze(-1), sy.to(dt)))


def _detector(seed=2, **cfg):
    from camouflage_multimodal_b200.fusion_model import build_multimodal_model
    torch.manual_seed(seed)
    det = build_multimodal_model(cfg)
    with torch.no_grad():
        for n, p in det.named_parameters():
            if "ln_" in n and n.endswith("weight"):
                p.uniform_(0.7, 1.3)
            elif n.endswith("bias"):
                p.normal_(0, 0.05)
    return det.cuda()


def _fusion_oracle(det, rows, node_ptr, kg, targets, heads):
    from oracle import fusion as ofus
    sd = {k: v.detach().cpu().double().requires_grad_(True) for k, v in det.state_dict().items()}
    rg = rows.detach().cpu().double().requires_grad_(True)
    kgd = kg.detach().cpu().double().requires_grad_(True)
    ptr = node_ptr.cpu().tolist()
    outs = []
    for g in range(len(ptr) - 1):
        m, i, e, s, _, _ = ofus.detector_forward(sd, rg[ptr[g]:ptr[g + 1]][None], kgd[None], heads=heads)
        outs.append((m, i, e, s))
    cat = [torch.cat([o[j] for o in outs]) for j in range(4)]
    loss = _fusion_loss(*cat, *targets)
    loss.backward()
    return cat, loss, {k: v.grad for k, v in sd.items()}, rg.grad, kgd.grad


@pytest.mark.parametrize("cfg", [dict(), dict(rg_dim=256, kg_dim=64, hidden_dim=256, num_heads=4, num_classes=3)])
def test_fusion_train_step_gradients_match_autograd_oracle(cfg):
    """Ragged batch of 3 images against the float64 oracle run per image: every parameter gradient plus the gradients
    with respect to the region embeddings and the (shared) KG rows.  Second config: Identity rg_proj, head_dim 64."""
    det = _detector(**cfg).train()
    det.dropout_p = 0.0
    gen = torch.Generator().manual_seed(9)
    sizes = [57, 131, 90]
    M, G, K = sum(sizes), 3, 13
    rows = (torch.randn((M, cfg.get("rg_dim", 128)), generator=gen).clamp_min(0) * 0.5).cuda().requires_grad_(True)
    kg = (torch.randn((K, cfg.get("kg_dim", 128)), generator=gen).clamp_min(0) * 0.25).cuda().requires_grad_(True)
    node_ptr = torch.tensor([0, 57, 188, 278], dtype=torch.int32, device="cuda")
    nc = cfg.get("num_classes", 2)
    targets = (torch.randint(0, nc, (G,), generator=gen), torch.randint(0, nc, (G,), generator=gen),
               torch.randint(0, 2, (G,), generator=gen).float(), torch.randint(0, 2, (G,), generator=gen).float())
    res = det.forward_ragged(rows, node_ptr, G, kg[None])
    outs = (res["mask_out"], res["instance_out"], res["edge_out"], res["score_out"])
    loss = _fusion_loss(*outs, *[t.cuda() for t in targets])
    loss.backward()
    ocat, oloss, og, org, okg = _fusion_oracle(det, rows, node_ptr, kg, targets, cfg.get("num_heads", 8))
    for got, ref in zip(outs, ocat):
        np.testing.assert_allclose(got.detach().cpu().numpy(), ref.detach().numpy(), rtol=0, atol=2e-4)
    assert abs(loss.item() - oloss.item()) <= 1e-4
    bad = {}
    named = dict(det.named_parameters())
    named["<rg rows>"], named["<kg rows>"] = rows, kg
    og["<rg rows>"], og["<kg rows>"] = org, okg
    for name, prm in named.items():
        ref = og[name].numpy()
        got = prm.grad.detach().cpu().numpy().astype(np.float64)
        assert got.shape == ref.shape, name
        tol = 3e-4 * np.abs(ref).max() + 1e-6
        err = np.abs(got - ref).max() / tol
        if err > 1.0:
            bad[name] = err
    assert not bad, f"gradient mismatch (error / tolerance): {bad}"


def test_fusion_train_dense_forward_signature_and_dropout():
    """The reference's call shape under model.train(): dense [B,N,d] / [B,K,d] inputs, 4 outputs, dropout on:
    finite gradients for every parameter, different masks give different losses, eval afterwards still works."""
    det = _detector(4).train()
    gen = torch.Generator().manual_seed(1)
    rg = torch.randn((2, 40, 128), generator=gen).clamp_min(0).cuda()
    kg = torch.randn((1, 13, 128), generator=gen).clamp_min(0).mul(0.25).cuda().expand(2, -1, -1)
    losses = []
    for _ in range(2):
        det.zero_grad()
        mask, inst, edge, score = det(rg, kg)
        assert tuple(mask.shape) == (2, 2) and tuple(score.shape) == (2, 1)
        loss = mask.square().mean() + score.mean() + inst.sum() * 0.1 + edge.sum() * 0.1
        loss.backward()
        losses.append(loss.item())
        assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in det.parameters())
    assert losses[0] != losses[1]
    det.eval()
    with torch.no_grad():
        assert torch.isfinite(det(rg, kg)[0]).all()


def test_joint_step_gradients_flow_from_fusion_into_the_gnn():
    """configs[3] in small: loss on the image-level heads only, so every GNN gradient arrives through
    cm_fusion_backward's d_rg -> cm_gnn_backward's d_node_emb.  Oracle: autograd through both float64 restatements."""
    from camouflage_multimodal_b200 import training
    from oracle import fusion as ofus, gnn as ognn
    data, y, iy, ey = _batch(2, first=66)
    gnn = _model(9).train()
    gnn.p_conv = gnn.p_fc = 0.0
    det = _detector(6).train()
    det.dropout_p = 0.0
    kg = synth.synthetic_kg(13).cuda()
    ptr = torch.zeros(3, dtype=torch.int32, device="cuda")
    ptr[1:] = torch.cumsum(torch.bincount(data.batch, minlength=2), 0).to(torch.int32)
    tg = (torch.tensor([1, 0]), torch.tensor([0, 1]), torch.tensor([1.0, 0.0]), torch.tensor([0.0, 1.0]))

    def loss_fn(node_outs, res):
        return _fusion_loss(res["mask_out"], res["instance_out"], res["edge_out"], res["score_out"], *[t.cuda() for t in tg])

    loss = training.joint_train_step(gnn, det, data, ptr, 2, kg[None], loss_fn, optimizer=None, max_norm=None)
    gsd = {k: v.detach().cpu().double().requires_grad_("running" not in k) for k, v in gnn.state_dict().items() if "num_batches" not in k}
    fsd = {k: v.detach().cpu().double().requires_grad_(True) for k, v in det.state_dict().items()}
    r = ognn.rg_gnn_forward(gsd, data.x.cpu().double(), data.edge_index.cpu(), data.edge_attr.cpu().double(), training_bn=True)
    p = ptr.cpu().tolist()
    outs = [ofus.detector_forward(fsd, r["node_emb"][p[g]:p[g + 1]][None], kg.cpu().double()[None]) for g in range(2)]
    oloss = _fusion_loss(*[torch.cat([o[j] for o in outs]) for j in range(4)], *tg)
    oloss.backward()
    assert abs(loss.item() - oloss.item()) <= 1e-4
    bad = {}
    for name, prm in list(gnn.named_parameters()) + list(det.named_parameters()):
        ref = (gsd[name] if name in gsd else fsd[name]).grad
        if any(name.startswith(h) for h in ("fc_mask", "fc_instance", "fc_edge")):   # node heads carry no loss here
            assert prm.grad.abs().max().item() == 0.0
            continue
        ref = ref.numpy()
        got = prm.grad.detach().cpu().numpy().astype(np.float64)
        tol = 5e-4 * np.abs(ref).max() + 1e-6
        if np.abs(got - ref).max() > tol:
            bad[name] = np.abs(got - ref).max() / tol
    assert not bad, f"gradient mismatch (error / tolerance): {bad}"


def test_fusion_train_step_with_all_nine_dropout_masks():
    """Every dropout site of the fusion (both attention-probability dropouts, both FFNs, the fusion MLP, four heads)
    with explicit keep masks, p = 0.3, against the float64 oracle given the same masks."""
    from camouflage_multimodal_b200 import training
    det = _detector(8).train()
    pd = det.dropout_p
    assert abs(pd - 0.3) < 1e-9
    gen = torch.Generator().manual_seed(13)
    sizes = [44, 71]
    M, G, K, D, H = sum(sizes), 2, 13, 256, 8
    rows = (torch.randn((M, 128), generator=gen).clamp_min(0) * 0.5).cuda().requires_grad_(True)
    kg = (torch.randn((K, 128), generator=gen).clamp_min(0) * 0.25).cuda()
    node_ptr = torch.tensor([0, 44, 115], dtype=torch.int32, device="cuda")
    shapes = [(M, H, K), (M, 2 * D), (H, K, M), (G * K, 2 * D), (G, D)] + [(G, D // 2)] * 4
    masks = [(torch.rand(sh, generator=gen) >= pd).to(torch.uint8).cuda() for sh in shapes]
    targets = (torch.tensor([1, 0]), torch.tensor([0, 1]), torch.tensor([1.0, 0.0]), torch.tensor([0.0, 1.0]))
    fused, *outs = training.fusion_train_forward(det, rows, node_ptr, G, kg[None], masks=masks)
    loss = _fusion_loss(*outs, *[t.cuda() for t in targets])
    loss.backward()
    # oracle, image by image, with the same masks sliced to the image
    from oracle import fusion as ofus
    sd = {k: v.detach().cpu().double().requires_grad_(True) for k, v in det.state_dict().items()}
    rg = rows.detach().cpu().double().requires_grad_(True)
    mk = [m.cpu() for m in masks]
    ptr = node_ptr.cpu().tolist()
    res = []
    for g in range(G):
        a, b = ptr[g], ptr[g + 1]
        drop = {"r2k": mk[0][a:b].permute(1, 0, 2)[None], "ffn_rg": mk[1][a:b][None], "k2r": mk[2][:, :, a:b][None],
                "ffn_kg": mk[3][g * K:(g + 1) * K][None], "fusion_layer": mk[4][g][None], "mask_head": mk[5][g][None],
                "instance_head": mk[6][g][None], "edge_head": mk[7][g][None], "score_head": mk[8][g][None]}
        res.append(ofus.detector_forward(sd, rg[a:b][None], kg.cpu().double()[None], drop=drop, pdrop=pd))
    ocat = [torch.cat([r[j] for r in res]) for j in range(4)]
    oloss = _fusion_loss(*ocat, *targets)
    oloss.backward()
    for got, ref in zip(outs, ocat):
        np.testing.assert_allclose(got.detach().cpu().numpy(), ref.detach().numpy(), rtol=0, atol=3e-4)
    bad = {}
    named = dict(det.named_parameters())
    named["<rg rows>"] = rows
    og = {k: v.grad for k, v in sd.items()}
    og["<rg rows>"] = rg.grad
    for name, prm in named.items():
        ref = og[name].numpy()
        got = prm.grad.detach().cpu().numpy().astype(np.float64)
        tol = 3e-4 * np.abs(ref).max() + 1e-6
        if np.abs(got - ref).max() > tol:
            bad[name] = np.abs(got - ref).max() / tol
    assert not bad, f"gradient mismatch (error / tolerance): {bad}"


def _grad_check(named, og, rel=3e-4):
    bad = {}
    for name, prm in named.items():
        ref = og[name].numpy()
        got = prm.grad.detach().cpu().numpy().astype(np.float64)
        tol = rel * np.abs(ref).max() + 1e-6
        if np.abs(got - ref).max() > tol:
            bad[name] = np.abs(got - ref).max() / tol
    assert not bad, f"gradient mismatch (error / tolerance): {bad}"


def test_late_fusion_detector_trains():
    """fusion_type='late' (configs/multimodal_config.yaml: fusion_type) under model.train(): forward with explicit
    dropout masks on both nn.Dropout sites + the four heads, gradients against float64 autograd over the oracle."""
    from camouflage_multimodal_b200 import training
    from oracle import fusion as ofus
    det = _detector(4, fusion_type="late").train()
    pd = det.dropout_p
    gen = torch.Generator().manual_seed(31)
    sizes = [37, 90, 5]
    M, G, K, D = sum(sizes), 3, 13, 256
    rows = (torch.randn((M, 128), generator=gen).clamp_min(0) * 0.5).cuda().requires_grad_(True)
    kg = (torch.randn((K, 128), generator=gen).clamp_min(0) * 0.25).cuda()
    node_ptr = torch.tensor([0, 37, 127, 132], dtype=torch.int32, device="cuda")
    shapes = [(G, D), (G, D // 2), None, None, None] + [(G, D // 4)] * 4
    masks = [None if sh is None else (torch.rand(sh, generator=gen) >= pd).to(torch.uint8).cuda() for sh in shapes]
    tg = (torch.tensor([1, 0, 1]), torch.tensor([0, 1, 1]), torch.tensor([1.0, 0.0, 1.0]), torch.tensor([0.0, 1.0, 0.5]))
    fused, *outs = training.fusion_train_forward(det, rows, node_ptr, G, kg[None], masks=masks)
    assert fused.shape == (G, D // 2)
    loss = _fusion_loss(*outs, *[t.cuda() for t in tg])
    loss.backward()
    sd = {k: v.detach().cpu().double().requires_grad_(True) for k, v in det.state_dict().items()}
    rg = rows.detach().cpu().double().requires_grad_(True)
    mk = [None if m is None else m.cpu() for m in masks]
    ptr = node_ptr.cpu().tolist()
    res = []
    for g in range(G):
        drop = {"late_0": mk[0][g][None], "late_1": mk[1][g][None], "mask_head": mk[5][g][None], "instance_head": mk[6][g][None],
                "edge_head": mk[7][g][None], "score_head": mk[8][g][None]}
        res.append(ofus.detector_forward(sd, rg[ptr[g]:ptr[g + 1]][None], kg.cpu().double()[None], fusion_type="late", drop=drop, pdrop=pd))
    ocat = [torch.cat([r[j] for r in res]) for j in range(4)]
    _fusion_loss(*ocat, *tg).backward()
    for got, ref in zip(outs, ocat):
        np.testing.assert_allclose(got.detach().cpu().numpy(), ref.detach().numpy(), rtol=0, atol=1e-4)
    named = dict(det.named_parameters())
    named["<rg rows>"] = rows
    og = {k: v.grad for k, v in sd.items()}
    og["<rg rows>"] = rg.grad
    _grad_check(named, og)
    # the module-level forward (reference signature) trains too and draws its own masks
    det.zero_grad()
    out = det(rows.detach().view(1, M, 128), kg[None])
    assert len(out) == 4 and out[0].requires_grad
    out[0].sum().backward()
    assert det.fusion.fusion[0].weight.grad.abs().sum() > 0


@pytest.mark.parametrize("kind", ["cross_attention", "late"])
def test_standalone_fusion_modules_train(kind):
    """CrossAttentionFusion / LateFusion on their own (fusion_model.py:16-171 are ordinary trainable modules): train-mode
    forward returns a differentiable ``fused``; dropout off -> gradients against float64 autograd over the oracle."""
    from camouflage_multimodal_b200.fusion_model import CrossAttentionFusion, LateFusion
    from oracle import fusion as ofus
    torch.manual_seed(12)
    mod = (CrossAttentionFusion(128, 128, 256, 8, dropout=0.0) if kind == "cross_attention" else LateFusion(128, 128, 256, dropout=0.0)).cuda().train()
    gen = torch.Generator().manual_seed(4)
    B, N, K = 2, 50, 13
    rg = (torch.randn((B, N, 128), generator=gen).clamp_min(0) * 0.5).cuda().requires_grad_(True)
    kg = (torch.randn((B, K, 128), generator=gen).clamp_min(0) * 0.25).cuda().requires_grad_(True)
    fused, attn = mod(rg, kg)
    w = torch.randn(fused.shape, generator=gen).cuda()
    (fused * w).sum().backward()
    sd = {"fusion." + k: v.detach().cpu().double().requires_grad_(True) for k, v in mod.state_dict().items()}
    rgo, kgo = rg.detach().cpu().double().requires_grad_(True), kg.detach().cpu().double().requires_grad_(True)
    if kind == "cross_attention":
        f_ref, a_ref = ofus.cross_attention_fusion(sd, rgo, kgo, 8)
        assert attn["rg2kg"].shape == (B, N, K) and not attn["rg2kg"].requires_grad
        np.testing.assert_allclose(attn["rg2kg"].cpu().numpy(), a_ref["rg2kg"].detach().numpy(), rtol=0, atol=1e-3)
    else:
        f_ref, _ = ofus.late_fusion(sd, rgo, kgo)
        assert attn is None
    (f_ref * w.cpu().double()).sum().backward()
    np.testing.assert_allclose(fused.detach().cpu().numpy(), f_ref.detach().numpy(), rtol=0, atol=2e-4)
    named = {"fusion." + k: v for k, v in mod.named_parameters()}
    named["<rg>"], named["<kg>"] = rg, kg
    og = {k: v.grad for k, v in sd.items()}
    og["<rg>"], og["<kg>"] = rgo.grad, kgo.grad
    _grad_check(named, og)
    # with dropout on, the forward still runs and differs from the eval output
    mod.dropout_p = 0.3
    f2, _ = mod(rg.detach(), kg.detach())
    mod.eval()
    with torch.no_grad():
        f3, _ = mod(rg.detach(), kg.detach())
    assert torch.isfinite(f2).all() and (f2 - f3).abs().max() > 1e-4


def test_asymmetric_edge_index_is_reported_without_stalling_the_backward():
    """The backward needs a structurally symmetric edge_index; the flag is read through pinned memory after the
    backward was queued (no device->host read inside the step): check_deferred_status(block=True) raises, and so does
    the next training forward; a symmetric graph leaves nothing pending."""
    from camouflage_multimodal_b200 import _lib, training
    from camouflage_multimodal_b200.data import Data
    m = _model(9).train()
    m.p_conv = m.p_fc = 0.0
    gen = torch.Generator().manual_seed(4)
    x = torch.randn((40, 15), generator=gen).cuda()
    src = torch.arange(0, 39)
    sym = torch.cat([torch.stack([src, src + 1]), torch.stack([src + 1, src])], 1)
    asym = torch.cat([sym, torch.tensor([[0], [17]])], 1)              # 0 -> 17 without 17 -> 0
    training.check_deferred_status(block=True)
    for ei, bad in ((sym, False), (asym, True)):
        data = Data(x=x, edge_index=ei.cuda(), edge_attr=torch.rand((ei.shape[1], 1), generator=gen).cuda())
        outs = m(data)
        sum(o.sum() for o in outs).backward()
        assert len(training._pending_status) == 1
        if bad:
            with pytest.raises(_lib.CmError, match="symmetric"):
                training.check_deferred_status(block=True)
        else:
            training.check_deferred_status(block=True)
        assert not training._pending_status
    # and through the next forward: one more asymmetric backward left pending, then a forward
    data = Data(x=x, edge_index=asym.cuda(), edge_attr=torch.rand((asym.shape[1], 1), generator=gen).cuda())
    sum(o.sum() for o in m(data)).backward()
    torch.cuda.synchronize()
    with pytest.raises(_lib.CmError, match="symmetric"):
        m(data)
    assert not training._pending_status


def _attention_path_grads():
    """One ragged training step of the default detector with every dropout mask fixed -> outputs + all gradients (CPU)."""
    from camouflage_multimodal_b200 import training
    det = _detector(6).train()
    gen = torch.Generator().manual_seed(19)
    sizes = [33, 410, 1, 257, 96]
    M, G, K = sum(sizes), len(sizes), 13
    rows = (torch.randn((M, 128), generator=gen).clamp_min(0) * 0.5).cuda().requires_grad_(True)
    kg = (torch.randn((K, 128), generator=gen).clamp_min(0) * 0.25).cuda().requires_grad_(True)
    node_ptr = torch.tensor([0] + list(np.cumsum(sizes)), dtype=torch.int32, device="cuda")
    shapes = [(M, 8, K), (M, 512), (8, K, M), (G * K, 512), (G, 256)] + [(G, 128)] * 4
    masks = [(torch.rand(sh, generator=gen) >= 0.1).to(torch.uint8).cuda() for sh in shapes]
    outs = training.fusion_train_forward(det, rows, node_ptr, G, kg[None], masks=masks)
    sum((o * torch.linspace(0.5, 1.5, o.numel(), device="cuda").view_as(o)).sum() for o in outs[1:]).backward()
    res = {"out%d" % i: o.detach().cpu() for i, o in enumerate(outs)}
    res.update({n: p.grad.detach().cpu() for n, p in det.named_parameters()})
    res["<rg rows>"], res["<kg rows>"] = rows.grad.cpu(), kg.grad.cpu()
    return res


def test_tiled_attention_kernels_agree_with_the_sequential_kernels(tmp_path):
    """head_dim 32, K <= 16 takes the tiled attention kernels of fusion_train.cu; CM_TRAIN_ATTN_SEQUENTIAL=1 (another
    process) keeps the one-row-at-a-time kernels.  Same sums in another order: every output and gradient agrees to
    1e-5 of its largest magnitude (graphs of 1 .. 410 rows, dropout masks on every site)."""
    import os
    import subprocess
    import sys
    here = _attention_path_grads()
    path = tmp_path / "sequential.pt"
    code = ("import sys, torch; sys.path.insert(0, %r); sys.path.insert(0, %r); import test_gpu_training as t; "
            "torch.save(t._attention_path_grads(), %r)") % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                           os.path.dirname(os.path.abspath(__file__)), str(path))
    subprocess.run([sys.executable, "-c", code], check=True, env=dict(os.environ, CM_TRAIN_ATTN_SEQUENTIAL="1"), timeout=600)
    there = torch.load(path)
    assert set(here) == set(there)
    for name in here:
        a, b = here[name].double(), there[name].double()
        assert torch.isfinite(a).all(), name
        assert float((a - b).abs().max()) <= 1e-5 * float(b.abs().max()) + 1e-12, name
