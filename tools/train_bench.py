"""BASELINE configs[3]: one training step (forward + backward through the region-graph message passing and the
cross-attention fusion, gradient all-reduce, clip, AdamW) at batch 16 per GPU, random-init weights, synthetic data.

  python tools/train_bench.py [--steps 20] [--warmup 5] [--batch 16]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/train_bench.py

Prints one JSON line (rank 0): images/s over all ranks, ms/step (CUDA events, max over ranks) and the per-kernel split.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=16)
    args = ap.parse_args()
    import torch.distributed as dist

    from camouflage_multimodal_b200 import _lib, synth, training
    from camouflage_multimodal_b200.fusion_model import build_multimodal_model
    from camouflage_multimodal_b200.region_graph import RegionGraphGNN, batch_to_data, build_region_graphs
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(0)
    gnn = RegionGraphGNN(15, 128, 2).to(dev).train()
    det = build_multimodal_model({}).to(dev).train()
    kg = synth.synthetic_kg(13).to(dev)
    if world > 1:
        for t in list(gnn.state_dict().values()) + list(det.state_dict().values()) + [kg]:
            dist.broadcast(t, src=0)
    B = args.batch
    b = synth.make_batch(rank * B, B, seed=1234, device=dev).as_decoded()
    params = list(gnn.parameters()) + list(det.parameters())
    from camouflage_multimodal_b200.train_loop import FusedAdamW, MultimodalLoss, RegionGraphLoss, region_targets
    # clip_grad_norm_(1.0) + AdamW(lr, weight_decay=1e-4) in cm_adamw_step (train.py:318, 356-357)
    opt = FusedAdamW(params, lr=1e-4, weight_decay=1e-4, max_norm=1.0)
    gen = torch.Generator().manual_seed(rank)
    img_y = torch.randint(0, 2, (B,), generator=gen).to(dev)
    edge_y = torch.randint(0, 2, (B,), generator=gen).float().to(dev)
    score_y = torch.rand((B,), generator=gen).to(dev)
    rg_loss, mm_loss = RegionGraphLoss(), MultimodalLoss()
    labels32 = b.labels.to(torch.int32)

    def step():
        # graph build (inference kernels) is part of the step: the reference builds graphs in its Dataset.__getitem__
        g = build_region_graphs(b.images, labels32, b.canny, label_cap=544)
        data = batch_to_data(g)
        n = data.x.shape[0]
        # per-region targets from the ground-truth masks (train.py:169-176); the synthetic set has one mask per image,
        # used for all three maps
        y, iy, ey = region_targets(g, labels32, b.gt, b.gt, b.gt)

        def loss_fn(node_outs, res):
            ln = rg_loss(node_outs[0], node_outs[1], node_outs[2], y[:n], iy[:n], ey[:n])
            lf = mm_loss(res["mask_out"], res["instance_out"], res["edge_out"], res["score_out"], img_y, edge_y, score_y)
            return ln + lf

        return training.joint_train_step(gnn, det, data, g.node_ptr, B, kg[None], loss_fn, optimizer=opt, max_norm=None), n

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        loss, n = step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    _lib.timing_enable(True)
    step()
    torch.cuda.synchronize()
    rep = _lib.timing_report()
    _lib.timing_enable(False)
    if rank == 0:
        tot = sum(v[1] for v in rep.values())
        print(json.dumps({"metric": "training images/sec (GNN + fusion forward+backward, grad all-reduce, clip, AdamW)",
                          "value": args.steps * B * world / (ms.item() / 1e3), "unit": "images/sec", "n_gpus": world,
                          "ms_per_step": ms.item() / args.steps, "batch_per_gpu": B, "nodes_per_step": n, "loss": float(loss.detach()),
                          "dtype": "f32", "kernel_ms_per_step": tot, "launches_per_step": sum(v[0] for v in rep.values()),
                          "kernels": {k: round(v[1], 4) for k, v in sorted(rep.items(), key=lambda kv: -kv[1][1])[:12]}}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
