"""CPU-side checks: the C-ABI library loads and exports every symbol include/cm_b200.h declares, the host
mirror keeps the reference's names / state-dict keys / error behaviour, nothing computes without a GPU, and
the N>1 sharding + all-reduce logic works over gloo."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_functions():
    src = open(os.path.join(ROOT, "include", "cm_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(cm_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from camouflage_multimodal_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "camouflage_multimodal_b200", "csrc"), "-j4"])
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = _header_functions()
    assert len(names) >= 19
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/cm_b200.h but not exported"
    assert set(names) == set(_lib.SIGNATURES), "ctypes binding and header drifted apart"
    assert _lib.load().cm_abi_version() == 1


def test_no_cpu_fallback():
    """Without a CUDA device every compute entry refuses loudly."""
    from camouflage_multimodal_b200 import _lib, metrics
    from camouflage_multimodal_b200.data import Data
    from camouflage_multimodal_b200.fusion_model import build_multimodal_model
    from camouflage_multimodal_b200.region_graph import RegionGraphGNN, build_region_graphs
    if torch.cuda.is_available():
        pytest.skip("CPU-only check")
    assert _lib.load().cm_device_check() == _lib.CM_ERR_DEVICE
    with pytest.raises(_lib.CmError):
        metrics.evaluate_segmentation(np.zeros((4, 4)), np.zeros((4, 4)))
    with pytest.raises(_lib.CmError):
        build_multimodal_model({}).eval()(torch.zeros(1, 3, 128), torch.zeros(1, 2, 128))
    with pytest.raises(_lib.CmError):
        RegionGraphGNN().eval()(Data(x=torch.zeros(2, 15), edge_index=torch.zeros((2, 0), dtype=torch.long),
                                     edge_attr=torch.zeros((0, 1))))
    with pytest.raises(_lib.CmError):
        build_region_graphs(torch.zeros(1, 4, 4, 3), torch.zeros(1, 4, 4, dtype=torch.int32),
                            torch.zeros(1, 4, 4, dtype=torch.uint8), label_cap=4)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "camouflage_multimodal_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
                assert "/root/reference" not in txt, f


def test_state_dict_keys_match_reference(golden_dir):
    from camouflage_multimodal_b200.fusion_model import build_multimodal_model
    from camouflage_multimodal_b200.region_graph import RegionGraphGNN
    z = np.load(os.path.join(golden_dir, "fusion.npz"))
    m = build_multimodal_model({})
    sd = m.state_dict()
    assert list(sd) == [str(k) for k in z["shapes_keys"]]
    for k, v in zip(z["shapes_keys"], z["shapes_vals"]):
        assert str(tuple(sd[str(k)].shape)) == str(v)
    late = build_multimodal_model({"fusion_type": "late"})
    assert list(late.state_dict()) == [str(k) for k in z["late_shapes_keys"]]
    zg = np.load(os.path.join(golden_dir, "region_graph.npz"))
    g = RegionGraphGNN(15, 128, 2).state_dict()
    assert set(g) == {str(k) for k in zg["gnn_shapes_keys"]}
    for k, v in zip(zg["gnn_shapes_keys"], zg["gnn_shapes_vals"]):
        assert str(tuple(g[str(k)].shape)) == str(v)


def test_reference_error_behaviour():
    from camouflage_multimodal_b200.fusion_model import CrossAttentionFusion, _collapse_to_3d, build_multimodal_model
    with pytest.raises(ValueError, match="Unknown fusion_type"):
        build_multimodal_model({"fusion_type": "mid"})
    with pytest.raises(ValueError, match="must be 2D/3D/4D tensor"):
        _collapse_to_3d(torch.zeros(1, 1, 1, 1, 1), "rg_embeddings")
    assert tuple(_collapse_to_3d(torch.zeros(1, 13, 1, 128), "kg").shape) == (1, 13, 128)
    assert tuple(_collapse_to_3d(torch.zeros(2, 1, 7, 128), "kg").shape) == (2, 7, 128)
    assert tuple(_collapse_to_3d(torch.zeros(2, 3, 5, 128), "kg").shape) == (2, 15, 128)
    f = CrossAttentionFusion(rg_dim=256, kg_dim=128, hidden_dim=256)
    assert isinstance(f.rg_proj, torch.nn.Identity) and isinstance(f.kg_proj, torch.nn.Linear)
    cfg = build_multimodal_model({"hidden_dim": 128, "num_heads": 4, "num_classes": 3})
    assert cfg.mask_head[3].out_features == 3 and cfg.fusion.cross_attn_rg2kg.num_heads == 4


def test_data_batch_standins():
    from camouflage_multimodal_b200.data import Batch, Data
    a = Data(x=torch.zeros(3, 15), edge_index=torch.tensor([[0, 1], [1, 0]]), edge_attr=torch.ones(2, 1))
    b = Data(x=torch.ones(2, 15), edge_index=torch.tensor([[0, 1], [1, 0]]), edge_attr=torch.ones(2, 1))
    bt = Batch.from_data_list([a, b])
    assert bt.batch.tolist() == [0, 0, 0, 1, 1] and bt.edge_index.tolist() == [[0, 1, 3, 4], [1, 0, 4, 3]]
    assert bt.num_graphs == 2 and bt.ptr.tolist() == [0, 3, 5]


def test_summary_from_partial_matches_numpy():
    from camouflage_multimodal_b200 import metrics as gm
    rng = np.random.default_rng(0)
    m = rng.random((9, 7))
    part = np.concatenate([[9.0], m.sum(0), (m * m).sum(0)])
    s = gm.summary_from_partial(part)
    for j, k in enumerate(gm.KEYS):
        assert abs(s[k] - m[:, j].mean()) < 1e-12 and abs(s[f"{k}_std"] - m[:, j].std()) < 1e-9


_WORKER = r"""
import os, sys
sys.path.insert(0, sys.argv[1])
import numpy as np, torch, torch.distributed as dist
from camouflage_multimodal_b200.pipeline import allreduce_partials, shard_indices, summarize
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%s" % sys.argv[2], rank=int(sys.argv[3]), world_size=2)
rank = dist.get_rank()
rng = np.random.default_rng(5)
m = rng.random((11, 7))                      # per-image metrics of the whole set
mine = shard_indices(11, rank, 2)
part = torch.tensor(np.concatenate([[len(mine)], m[mine].sum(0), (m[mine] ** 2).sum(0)]), dtype=torch.float64)
conf = torch.tensor([rank + 1, 2, 3, 4], dtype=torch.int64)
part, conf = allreduce_partials(part, conf)
s = summarize(part)
assert part[0].item() == 11 and conf.tolist() == [3, 4, 6, 8]
keys = ("iou", "dice", "precision", "recall", "f1", "mae", "accuracy")
for j, k in enumerate(keys):
    assert abs(s[k] - m[:, j].mean()) < 1e-12 and abs(s[k + "_std"] - m[:, j].std()) < 1e-9
dist.destroy_process_group()
print("ok", rank)
"""


def test_sharded_metric_allreduce_gloo(tmp_path):
    """world_size-2 sharding: i mod W partition covers the set once; summed partials reproduce the
    single-process batch statistics."""
    from camouflage_multimodal_b200.pipeline import shard_indices
    assert sorted(shard_indices(11, 0, 2) + shard_indices(11, 1, 2)) == list(range(11))
    script = tmp_path / "w.py"
    script.write_text(_WORKER)
    port = str(29500 + os.getpid() % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, port, str(r)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT) for r in range(2)]
    for p in procs:
        out, _ = p.communicate(timeout=240)
        assert p.returncode == 0, out.decode()


_GRAD_WORKER = r"""
import sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
rank = int(sys.argv[3])
dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{sys.argv[2]}", rank=rank, world_size=2)
from camouflage_multimodal_b200.training import allreduce_gradients
ps = [torch.nn.Parameter(torch.zeros(3, 5)), torch.nn.Parameter(torch.zeros(7)), torch.nn.Parameter(torch.zeros(2))]
ps[0].grad = torch.full((3, 5), float(rank + 1))
ps[1].grad = torch.arange(7.0) * (rank + 1)
allreduce_gradients(ps)                      # ps[2] has no grad: skipped on both ranks
assert torch.equal(ps[0].grad, torch.full((3, 5), 3.0)) and torch.equal(ps[1].grad, torch.arange(7.0) * 3) and ps[2].grad is None
allreduce_gradients(ps, average=True)
assert torch.equal(ps[0].grad, torch.full((3, 5), 3.0))
dist.destroy_process_group()
"""


def test_gradient_allreduce_gloo(tmp_path):
    """The data-parallel training step's only collective: one flat SUM all-reduce of every gradient."""
    script = tmp_path / "g.py"
    script.write_text(_GRAD_WORKER)
    port = str(31500 + os.getpid() % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, port, str(r)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT) for r in range(2)]
    for p in procs:
        out, _ = p.communicate(timeout=240)
        assert p.returncode == 0, out.decode()


_BN_WORKER = r"""
import sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
rank = int(sys.argv[3])
dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{sys.argv[2]}", rank=rank, world_size=2)
from camouflage_multimodal_b200.training import allreduce_bn_sums
from camouflage_multimodal_b200.pipeline import allreduce_partials
torch.manual_seed(0)
x = torch.randn(37, 16, dtype=torch.float64) * 2 + 1          # pre-BatchNorm rows of the whole batch
mine = x[:15] if rank == 0 else x[15:]                        # uneven shards
sums = torch.cat([mine.sum(0), (mine * mine).sum(0), torch.tensor([float(mine.shape[0])], dtype=torch.float64)])
allreduce_bn_sums(sums)
n = sums[-1]
mean, var = sums[:16] / n, sums[16:32] / n - (sums[:16] / n) ** 2
bn = torch.nn.BatchNorm1d(16, momentum=1.0).double().train()  # momentum 1: running stats = this batch's
bn(x)
assert n.item() == 37 and torch.allclose(mean, bn.running_mean, atol=1e-12)
assert torch.allclose(var * n / (n - 1), bn.running_var, atol=1e-12)
# the S/E/F partial vector rides the same collective as the 15-element one
part, sod = torch.full((15,), float(rank + 1), dtype=torch.float64), torch.arange(528, dtype=torch.float64) * (rank + 1)
part, conf, sod = allreduce_partials(part, None, sod)
assert part.tolist() == [3.0] * 15 and torch.equal(sod, torch.arange(528, dtype=torch.float64) * 3) and conf is None
dist.destroy_process_group()
"""


def test_batchnorm_sums_allreduce_gloo(tmp_path):
    """Data-parallel BatchNorm: all-reduced (sum x, sum x^2, n) of uneven shards give the statistics
    nn.BatchNorm1d takes over the whole batch; the GPU side of the hook is tests/test_gpu_dist_training.py."""
    script = tmp_path / "b.py"
    script.write_text(_BN_WORKER)
    port = str(33500 + os.getpid() % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, port, str(r)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT) for r in range(2)]
    for p in procs:
        out, _ = p.communicate(timeout=240)
        assert p.returncode == 0, out.decode()


def test_embedding_matcher_matches_reference_class_golden():
    """tests/golden/embedding_matcher.json holds the outputs of the reference's own EmbeddingMatcher
    (tools/gen_golden_matcher.py); the mirror must reproduce every category choice, id list and tensor."""
    import json
    import sys
    import numpy as np
    import torch
    from camouflage_multimodal_b200.embedding_cache import EmbeddingMatcher
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tools"))
    import gen_golden_matcher as gg      # only its seeded input builder is used (no reference import at module level ...)
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "embedding_matcher.json")))
    rg, kg = gg.inputs()
    for tag, mapping in (("no_mapping", None), ("mapping", gold["mapping"])):
        m = EmbeddingMatcher(rg, kg, category_mapping=mapping)
        case = gold["cases"][tag]
        for n in gold["names"]:
            assert m.extract_category_from_filename(n) == case["category_of"][n]
            e, ids = m.get_kg_embedding_for_image(n, use_all_categories=False)
            assert ids == case["single"][n]["ids"]
            np.testing.assert_allclose(e.numpy(), np.array(case["single"][n]["kg"], dtype=np.float32), rtol=0, atol=0)
        e, ids = m.get_kg_embedding_for_image(gold["names"][0], use_all_categories=True)
        assert ids == case["all"]["ids"]
        np.testing.assert_array_equal(e.numpy(), np.array(case["all"]["kg"], dtype=np.float32))
        for flag in (True, False):
            md = m.create_matched_dataset(use_all_kg_categories=flag)
            ref = case[f"matched_{flag}"]
            assert len(md) == len(ref)
            for r, g in zip(md, ref):
                assert sorted(r.keys()) == g["keys"]
                assert (r["image_name"], r["category_ids"], r["num_rg_nodes"], r["num_kg_categories"]) == \
                       (g["image_name"], g["category_ids"], g["num_rg_nodes"], g["num_kg_categories"])
                assert abs(float(r["kg_embeddings"].double().sum()) - g["kg_sum"]) < 1e-12
                assert abs(float(r["rg_node_embeddings"].double().sum()) - g["rg_sum"]) < 1e-12


def test_embedding_cache_files_round_trip(tmp_path):
    import json
    import torch
    from camouflage_multimodal_b200.embedding_cache import EmbeddingMatcher, save_embedding_cache
    gen = torch.Generator().manual_seed(5)
    allemb = {f"COD10K-CAM-1-Aquatic-1-Batfish-{i}.jpg": {"node_embeddings": torch.randn((4 + i, 128), generator=gen),
                                                          "graph_embedding": torch.randn((1, 128), generator=gen), "num_nodes": 4 + i}
              for i in range(3)}
    summary = save_embedding_cache(allemb, str(tmp_path))
    assert summary["total_images"] == 3 and summary["embedding_dim"] == 128
    loaded = torch.load(tmp_path / "all_rg_embeddings.pt", weights_only=False)
    assert list(loaded) == list(allemb) and set(loaded[next(iter(loaded))]) == {"node_embeddings", "graph_embedding", "num_nodes"}
    one = torch.load(tmp_path / "COD10K-CAM-1-Aquatic-1-Batfish-1_embedding.pt", weights_only=False)
    assert one["num_nodes"] == 5 and one["embedding_dim"] == 128 and one["image_name"].endswith("Batfish-1.jpg")
    assert json.load(open(tmp_path / "embedding_summary.json"))["images"][one["image_name"]]["node_embedding_shape"] == [5, 128]
    kg = {"Fish": torch.randn(128, generator=gen), "Bird": torch.randn(128, generator=gen)}
    torch.save(kg, tmp_path / "kg.pt")
    m = EmbeddingMatcher(str(tmp_path / "all_rg_embeddings.pt"), str(tmp_path / "kg.pt"))
    rows, ptr, kgb = m.matched_ragged_batch(list(allemb), device="cpu")
    assert rows.shape == (4 + 5 + 6, 128) and ptr.tolist() == [0, 4, 9, 15] and kgb.shape == (1, 2, 128)


def test_pack_labels10_bit_layout():
    """Label i occupies bits [10 i, 10 i + 10) of a little-endian stream (what cm_unpack_labels10 reads)."""
    import numpy as np
    import torch
    from camouflage_multimodal_b200.pipeline import pack_labels10
    lab = np.array([[[1, 2, 1023, 0, 513, 7, 64, 999, 5, 6, 7, 8, 9, 10, 11, 12]]], dtype=np.int64)
    stream = pack_labels10(torch.as_tensor(lab)).numpy().reshape(-1)
    assert stream.shape == (20,)
    big = int.from_bytes(stream.tobytes(), "little")
    assert [(big >> (10 * i)) & 1023 for i in range(16)] == lab.reshape(-1).tolist()


def test_keep_masks_one_draw_per_probability():
    """training.keep_masks: uint8 views of one byte buffer per distinct p, 16-byte aligned starts, None where the site
    has no dropout; the kept fraction is 1 - p."""
    from camouflage_multimodal_b200.training import keep_masks
    torch.manual_seed(0)
    shapes = [(1000, 33), None, (7, 5, 11), (64,), (300, 128)]
    probs = [0.3, 0.3, 0.3, 0.0, 0.2]
    masks = keep_masks(shapes, probs, "cpu")
    assert masks[1] is None and masks[3] is None
    for i in (0, 2, 4):
        m = masks[i]
        assert m.dtype == torch.uint8 and tuple(m.shape) == shapes[i] and m.is_contiguous() and m.data_ptr() % 16 == 0
        assert set(m.unique().tolist()) <= {0, 1}
    assert abs(masks[0].float().mean().item() - 0.7) < 0.01 and abs(masks[4].float().mean().item() - 0.8) < 0.01
    assert masks[0].untyped_storage().data_ptr() == masks[2].untyped_storage().data_ptr()      # same draw
    assert masks[0].untyped_storage().data_ptr() != masks[4].untyped_storage().data_ptr()
    assert keep_masks([(4, 4)], [0.0], "cpu") == [None]


_ISOLATION_WORKER = r"""
import os, sys, subprocess, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
if "--train-only" in sys.argv:      # what bench.py's child does with its environment: a process group of the children's own
    dist.init_process_group("gloo")
    t = torch.ones(1) * (int(os.environ["RANK"]) + 1)
    dist.all_reduce(t)
    if dist.get_rank() == 0:
        print('{"value": %d}' % int(t.item()), flush=True)
    dist.destroy_process_group()
    sys.exit(0)
import argparse, bench
bench.__file__ = os.path.abspath(__file__)          # train_block_isolated re-runs "this file" with --train-only
args = argparse.Namespace(steps=2, warmup=3, train_batch=16, train_all_kernels=False, train_inline_build=False,
                          train_host_profile=False, train_timeout=120)
world, rank = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"])
res = bench.train_block_isolated(args, world, rank)
assert (res == {"value": 3}) if rank == 0 else (res is None), res
dist.init_process_group("gloo")                     # the parent's own group afterwards, on the launcher's store
t = torch.ones(1) * 10
dist.all_reduce(t)
assert int(t.item()) == 10 * world
dist.destroy_process_group()
print("rank", rank, "ok", flush=True)
"""


def test_bench_training_block_isolation_under_torchrun(tmp_path):
    """bench.train_block_isolated: every rank starts a child with RANK / WORLD_SIZE kept, MASTER_PORT + 23 and no agent
    store, the children rendezvous among themselves, rank 0's child's JSON line comes back, and the parents' own process
    group works afterwards (2 ranks under torch.distributed.run, gloo).  Failure paths: a child that exits non-zero and a
    child that outlives the limit both yield {"unavailable": ...}."""
    script = tmp_path / "iso.py"
    script.write_text(_ISOLATION_WORKER)
    port = str(31000 + os.getpid() % 2000)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", port, str(script), ROOT], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=300)
    out = r.stdout.decode()
    assert r.returncode == 0, out
    assert "rank 0 ok" in out and "rank 1 ok" in out
    import argparse
    sys.path.insert(0, ROOT)
    import bench
    args = argparse.Namespace(steps=2, warmup=3, train_batch=16, train_all_kernels=False, train_inline_build=False,
                              train_host_profile=False, train_timeout=1)
    assert "did not finish" in bench.train_block_isolated(args, 1, 0)["unavailable"]      # import torch alone takes longer than 1 s
    if not torch.cuda.is_available():
        args.train_timeout = 120
        assert "exited with code" in bench.train_block_isolated(args, 1, 0)["unavailable"]    # no CUDA device: SystemExit in the child


def test_running_statistics_update_matches_batchnorm1d():
    """training.update_bn_running_stats (the _foreach form) == what nn.BatchNorm1d does with the same batch in train
    mode: momentum 0.1, unbiased variance, num_batches_tracked; Python-int and 0-dim-tensor row counts agree."""
    import torch.nn as nn
    from camouflage_multimodal_b200.training import update_bn_running_stats
    torch.manual_seed(3)
    n, h = 37, 16

    class M:
        pass
    xs = [torch.randn(n, h) * (k + 1) + k for k in range(4)]
    ref = [nn.BatchNorm1d(h).train() for _ in range(4)]
    for bn, x in zip(ref, xs):
        bn(x)
        bn(x)
    outs = []
    for count in (n, torch.tensor(float(n))):
        m = M()
        m.bn1, m.bn2, m.bn3, m.bn4 = (nn.BatchNorm1d(h) for _ in range(4))
        mean = torch.stack([x.mean(0) for x in xs])
        var = torch.stack([x.var(0, unbiased=False) for x in xs])
        update_bn_running_stats(m, mean, var, count)
        update_bn_running_stats(m, mean, var, count)
        outs.append(m)
        for bn, r in zip((m.bn1, m.bn2, m.bn3, m.bn4), ref):
            torch.testing.assert_close(bn.running_mean, r.running_mean, rtol=1e-6, atol=1e-6)
            torch.testing.assert_close(bn.running_var, r.running_var, rtol=1e-5, atol=1e-6)
            assert int(bn.num_batches_tracked) == 2
    for a, b in zip((outs[0].bn1, outs[0].bn4), (outs[1].bn1, outs[1].bn4)):
        torch.testing.assert_close(a.running_var, b.running_var, rtol=1e-6, atol=1e-7)
    m = outs[0]
    m.bn2.momentum = 0.3                      # mixed momenta: the per-layer loop
    before = m.bn2.running_mean.clone()
    update_bn_running_stats(m, mean, var, n)
    torch.testing.assert_close(m.bn2.running_mean, before * 0.7 + mean[1] * 0.3, rtol=1e-6, atol=1e-6)
