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
