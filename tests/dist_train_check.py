"""Worker for tests/test_gpu_dist_training.py (not collected by pytest): data-parallel GNN training step with
BatchNormSync + gradient all-reduce on `world` ranks == the single-process step on the whole batch.

  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port P tests/dist_train_check.py [--backend nccl|gloo]

Each rank takes images i with i mod world == rank of one synthetic batch; the loss is a SUM over nodes (additive over
ranks), so after the gradient SUM-all-reduce every rank must hold the single-process gradients.  With --backend gloo all
ranks may share cuda:0 (the 1-GPU box)."""
import argparse
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="nccl")
    ap.add_argument("--images", type=int, default=4)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    ndev = torch.cuda.device_count()
    dev = torch.device("cuda", local % ndev)
    torch.cuda.set_device(dev)
    dist.init_process_group(args.backend)
    from camouflage_multimodal_b200 import synth, training
    from camouflage_multimodal_b200.region_graph import RegionGraphGNN, batch_to_data, build_region_graphs

    def model():
        torch.manual_seed(3)
        m = RegionGraphGNN(15, 128, 2)
        with torch.no_grad():
            for bn in (m.bn1, m.bn2, m.bn3, m.bn4):
                bn.weight.uniform_(0.5, 1.5)
                bn.bias.normal_(0, 0.2)
        m.p_conv = m.p_fc = 0.0                      # dropout off: masks are random draws
        return m.to(dev).train()

    B = args.images
    hb = synth.make_batch(0, B, seed=77, n_regions=[90 + 17 * i for i in range(B)])

    def data_of(idx):
        d = hb.to(dev)
        g = build_region_graphs(d.images[idx], d.labels[idx], d.canny[idx], label_cap=256)
        return batch_to_data(g)

    def loss_of(outs, gen_seed):
        mask, inst, edge, emb = outs
        return (mask[:, 1].sigmoid().sum() + 0.5 * inst[:, 0].pow(2).sum() + edge.tanh().sum() + 0.01 * emb.pow(2).sum())

    # ---- data parallel: this rank's shard, BatchNorm statistics and gradients all-reduced
    mine = list(range(rank, B, world))
    m_dp = model()
    with training.BatchNormSync(dev, 128) as sync:
        outs = training.gnn_train_forward(m_dp, data_of(mine))
        loss = loss_of(outs, 0)
        loss.backward()
        n_calls = sync.calls
    params = list(m_dp.parameters())
    training.allreduce_gradients(params)
    tot = loss.detach().clone()
    dist.all_reduce(tot)
    # ---- single process: the whole batch, rank order = shard concatenation (BatchNorm / sums are order-free)
    m_sp = model()
    order = [i for r in range(world) for i in range(r, B, world)]
    outs = training.gnn_train_forward(m_sp, data_of(order))
    loss_sp = loss_of(outs, 0)
    loss_sp.backward()
    bad = {}
    for (name, a), b in zip(m_dp.named_parameters(), m_sp.parameters()):
        ref = b.grad.double().cpu().numpy()
        got = a.grad.double().cpu().numpy()
        if name.startswith("conv") and name.endswith(".bias"):
            # a bias in front of BatchNorm has an identically zero gradient (the batch mean removes it): both are rounding noise
            if max(np.abs(ref).max(), np.abs(got).max()) > 1e-3:
                bad[name] = float(np.abs(got).max())
            continue
        tol = 2e-4 * np.abs(ref).max() + 1e-6
        if np.abs(got - ref).max() > tol:
            bad[name] = float(np.abs(got - ref).max() / tol)
    for name in ("bn1", "bn4"):
        a, b = getattr(m_dp, name), getattr(m_sp, name)
        if (a.running_mean - b.running_mean).abs().max() > 1e-5 or (a.running_var - b.running_var).abs().max() > 1e-5:
            bad[name + ".running"] = float((a.running_var - b.running_var).abs().max())
    ok = (not bad) and abs(float(tot) - float(loss_sp)) <= 2e-4 * abs(float(loss_sp)) and n_calls == 8
    # without the hook the shard statistics differ: the check must be able to fail
    m_loc = model()
    o2 = training.gnn_train_forward(m_loc, data_of(mine))
    differs = world == 1 or float((o2[0] - training.gnn_train_forward(model(), data_of(order))[0][: o2[0].shape[0]]).abs().max()) > 1e-4
    flag = torch.tensor([1.0 if ok and differs else 0.0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(("DIST_TRAIN_OK" if flag.item() == 1.0 else "DIST_TRAIN_FAIL"), "world", world, "calls", n_calls, "loss", float(tot),
              float(loss_sp), "bad", bad, "differs", differs, flush=True)
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1.0 else 1)


if __name__ == "__main__":
    main()
