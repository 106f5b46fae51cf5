"""GPU parity: paint + metric reductions (A8, A9, A13) and the whole batched pipeline vs the oracle.
Confusion counts bit-exact; per-image ratio metrics bit-exact (float32 arithmetic restated); MAE <= 1e-9
given the same probabilities; end to end |dMAE| <= 1e-3 and mask logits <= 1e-2 (BASELINE.json)."""
import os

import numpy as np
import pytest
import torch

from camouflage_multimodal_b200 import synth

pytestmark = pytest.mark.gpu

from test_oracle_golden import fusion_state_dict, gnn_state_dict  # noqa: E402


def test_metrics_match_reference_golden(golden_dir):
    from camouflage_multimodal_b200 import metrics as gm
    from oracle import metrics as om
    z = np.load(os.path.join(golden_dir, "metrics.npz"))
    preds, gts = z["preds"], z["gts"]
    for i in range(len(preds)):
        # the device API computes on float32 arrays; compare with the oracle on the same rounded inputs
        p32, g32 = preds[i].astype(np.float32), gts[i].astype(np.float32)
        ref = om.evaluate_segmentation(p32, g32)
        got = gm.evaluate_segmentation(preds[i], gts[i])
        assert set(got) == set(om.KEYS)
        for k in ("iou", "dice", "precision", "recall", "f1", "accuracy"):
            assert got[k] == ref[k], (i, k)                          # bit-exact
        mae64 = float(np.abs(p32.astype(np.float64) - g32.astype(np.float64)).mean())
        assert abs(got["mae"] - mae64) < 1e-9
        # and against the reference's float64-input numbers themselves
        for j, k in enumerate(om.KEYS):
            assert abs(got[k] - z["per_image"][i][j]) < 1e-6, (i, k)
    be = gm.batch_evaluate(list(preds), list(gts))
    for k, v in zip(z["batch_keys"], z["batch_vals"]):
        assert abs(be[str(k)] - float(v)) < 1e-6, k
    assert gm.calculate_iou(preds[0], gts[0]) == gm.evaluate_segmentation(preds[0], gts[0])["iou"]
    p, r, f = gm.calculate_precision_recall_f1(preds[0], gts[0])
    assert 0 <= p <= 1 and 0 <= r <= 1 and 0 <= f <= 1
    assert abs(gm.calculate_mae(preds[0], gts[0]) - float(np.abs(preds[0] - gts[0]).mean())) < 1e-6


def test_class_confusion():
    from camouflage_multimodal_b200 import metrics as gm
    from oracle import metrics as om
    g = torch.Generator().manual_seed(0)
    logits = torch.randn(1000, 2, generator=g)
    true = torch.randint(0, 2, (1000,), generator=g)
    c = gm.class_confusion(logits.cuda(), true.cuda()).cpu().tolist()
    ref = om.image_level_f1(logits.argmax(1).numpy(), true.numpy())
    assert tuple(c) == ref["counts"]
    got = gm.f1_from_confusion(*c)
    assert abs(got["f1_avg"] - ref["f1_avg"]) < 1e-12


@pytest.fixture(scope="module")
def models(golden_dir):
    from camouflage_multimodal_b200.fusion_model import build_multimodal_model
    from camouflage_multimodal_b200.region_graph import RegionGraphGNN
    zg = np.load(os.path.join(golden_dir, "region_graph.npz"))
    zf = np.load(os.path.join(golden_dir, "fusion.npz"))
    gnn = RegionGraphGNN(15, 128, 2)
    gnn.load_state_dict(gnn_state_dict(zg))
    det = build_multimodal_model({})
    det.load_state_dict(fusion_state_dict(zf))
    kg = torch.from_numpy(np.load(os.path.join(golden_dir, "kg_embeddings.npz"))["kg"]).reshape(13, 128)
    return gnn.cuda().eval(), det.cuda().eval(), kg.cuda(), gnn_state_dict(zg), fusion_state_dict(zf)


@pytest.mark.parametrize("fix", [False, True])
def test_pipeline_vs_oracle(models, fix):
    from camouflage_multimodal_b200.pipeline import EvalPipeline
    from oracle import fusion as ofus, gnn as ognn, metrics as om, region_graph as org
    gnn, det, kg, gsd, fsd = models
    B = 3
    hb = synth.make_batch(200, B, seed=1234)
    d = hb.to("cuda")
    pipe = EvalPipeline(gnn, det, kg, label_cap=544, label_offset_fix=fix, want_pred_mask=True)
    out = pipe.run(d.images, d.labels, d.canny, d.gt)
    torch.cuda.synchronize()
    per = out["per_image"].cpu().numpy()
    ptr = out["graph"].node_ptr.cpu().numpy()
    for i in range(B):
        img, lab, can = hb.images[i].numpy(), hb.labels[i].numpy(), hb.canny[i].numpy()
        x, ei, w, id_of = org.build_region_graph(img.astype(np.float64), lab.astype(np.int64), can.astype(bool))
        with torch.no_grad():
            r = ognn.rg_gnn_forward(gsd, torch.from_numpy(x), torch.from_numpy(ei), torch.from_numpy(w))
            mask, inst, edge, score, _, _ = ofus.detector_forward(fsd, r["node_emb"][None], kg.cpu()[None])
        n0, n1 = ptr[i], ptr[i + 1]
        np.testing.assert_allclose(out["node_mask_out"][n0:n1].cpu().numpy(), r["mask_out"].numpy(), rtol=0, atol=1e-2)
        np.testing.assert_allclose(out["mask_out"][i].cpu().numpy(), mask[0].numpy(), rtol=0, atol=1e-2)
        np.testing.assert_allclose(out["score_out"][i].cpu().numpy(), score[0].numpy(), rtol=0, atol=1e-2)
        gt = hb.gt[i].numpy() / 255.0
        # (1) the reductions alone: oracle paint of the DEVICE logits -> counts must be bit-exact
        pm_dev = om.paint_mask(out["node_mask_out"][n0:n1].cpu().numpy(), lab, label_offset_fix=fix, id_of_label=id_of)
        np.testing.assert_allclose(out["pred_mask"][i].cpu().numpy(), pm_dev, rtol=0, atol=1e-6)
        pm32 = out["pred_mask"][i].cpu().numpy().astype(np.float64)
        tp, sp, sg, neq, n = om.confusion_counts(pm32, gt)
        assert out["counts"][i].cpu().tolist() == [tp, sp, sg, neq]
        ref = om.metrics_from_counts(tp, sp, sg, neq, n, float(np.abs(pm32 - gt).sum()))
        for j, k in enumerate(om.KEYS):
            assert abs(per[i, j] - ref[k]) < (1e-9 if k == "mae" else 1e-15), k
        # (2) end to end against the all-oracle chain: BASELINE tolerance on MAE
        e2e = om.evaluate_segmentation(om.paint_mask(r["mask_out"].numpy(), lab, label_offset_fix=fix, id_of_label=id_of), gt)
        assert abs(per[i, 5] - e2e["mae"]) <= 1e-3
    # batch statistics from the partial-sum vector
    from camouflage_multimodal_b200.pipeline import summarize
    s = summarize(out["partial"])
    for j, k in enumerate(om.KEYS):
        assert abs(s[k] - per[:, j].mean()) < 1e-12 and abs(s[f"{k}_std"] - per[:, j].std()) < 1e-9


def test_pipeline_batch32_properties(models):
    """Config-2 batch size: idempotence across calls and buffer reuse, counts add up, probabilities in [0,1]."""
    from camouflage_multimodal_b200.pipeline import EvalPipeline, HostStager
    gnn, det, kg, _, _ = models
    B = 32
    hb = synth.make_batch(300, B, seed=1234)
    pipe = EvalPipeline(gnn, det, kg, label_cap=544)
    st = HostStager(B, 256, 256, torch.device("cuda"))
    st.fill(hb)
    out = st.step(pipe)
    torch.cuda.synchronize()
    first = {k: out[k].clone() for k in ("counts", "per_image", "mask_out", "partial")}
    out = st.step(pipe)
    torch.cuda.synchronize()
    assert torch.equal(first["counts"], out["counts"])
    assert torch.allclose(first["per_image"], out["per_image"], atol=1e-12)
    assert torch.equal(first["mask_out"], out["mask_out"])
    assert torch.equal(st.h_mask_out, out["mask_out"].cpu())
    c = out["counts"].cpu().numpy()
    assert (c[:, 0] <= np.minimum(c[:, 1], c[:, 2])).all() and (c[:, 3] <= 65536).all()
    gt_pos = (hb.gt.numpy() > 127).reshape(B, -1).sum(1)
    assert np.array_equal(c[:, 2], gt_pos)
    p = out["mask_prob"][: int(out["graph"].node_ptr[-1])]
    assert (p >= 0).all() and (p <= 1).all()
    assert int(out["graph"].status.item()) == 0


def test_host_pipeline_graph_replay_equals_eager(models):
    """The captured step (CUDA graph, H2D on a copy stream, rotating slots) returns exactly what the eagerly
    launched kernels return, batch after batch, including when a slot is reused."""
    from camouflage_multimodal_b200.pipeline import EvalPipeline, HostPipeline
    gnn, det, kg, _, _ = models
    B = 4
    dev = torch.device("cuda")
    pipe = EvalPipeline(gnn, det, kg, label_cap=544)
    hp = HostPipeline(pipe, B, 256, 256, dev, slots=2, use_graph=True)
    batches = [synth.make_batch(50 + 7 * j, B, seed=1234) for j in range(5)]
    pinned = [tuple(t.pin_memory() for t in (b.images, b.labels, b.canny, b.gt)) for b in batches]
    got = []
    for j, hb in enumerate(pinned):
        st = hp.submit(j, *hb)
        st.ev_done.synchronize()
        got.append((st.h_mask_out.clone(), st.h_per_image.clone()))
    eager = EvalPipeline(gnn, det, kg, label_cap=544)
    for j, b in enumerate(batches):
        d = b.to("cuda")
        out = eager.run(d.images, d.labels, d.canny, d.gt)
        assert torch.equal(got[j][0], out["mask_out"].cpu())
        assert torch.equal(got[j][1], out["per_image"].cpu())


def test_lanes_and_decoded_inputs_equal_float_path(models):
    """Two lanes (disjoint scratch, own streams, captured steps) fed uint8 images + int16 label maps return what one
    eager pipeline returns for the same values as float32 images + int32 labels: metrics to 1e-6, logits to 1e-3
    (the two feature kernels agree to ~1e-7; the fp16 tensor-core operands of the dense layers turn a last-bit
    feature difference into ~1e-4 on logits of magnitude 4 -- the stated tolerance is 1e-2)."""
    from camouflage_multimodal_b200.pipeline import EvalPipeline, HostPipeline
    gnn, det, kg, _, _ = models
    B = 4
    dev = torch.device("cuda")
    pipes = [EvalPipeline(gnn, det, kg, label_cap=544, lane=k) for k in range(2)]
    hp = HostPipeline(pipes, B, 256, 256, dev, slots=4, use_graph=True, image_dtype=torch.uint8, label_dtype=torch.int16)
    batches = [synth.make_batch(90 + 5 * j, B, seed=1234).as_decoded() for j in range(6)]
    pinned = [tuple(t.pin_memory() for t in (b.images, b.labels, b.canny, b.gt)) for b in batches]
    slots = [hp.submit(j, *hb) for j, hb in enumerate(pinned[:4])]       # four batches in flight on two lanes
    hp.synchronize()
    got = [(st.h_mask_out.clone(), st.h_per_image.clone()) for st in slots]
    for j in (4, 5):                                                      # slot reuse; batch 5 through the one-copy staging layout
        if j == 5:
            hb = hp.host_batch().fill(batches[j].images, batches[j].labels, batches[j].canny, batches[j].gt)
            assert hb.flat.is_pinned() and hb.payload_bytes == hp.h2d_bytes
            st = hp.submit(j, hb)
        else:
            st = hp.submit(j, *pinned[j])
        st.ev_done.synchronize()
        got.append((st.h_mask_out.clone(), st.h_per_image.clone()))
    eager = EvalPipeline(gnn, det, kg, label_cap=544, lane=7)
    for j, b in enumerate(batches):
        d = b.as_float().to("cuda")
        out = eager.run(d.images, d.labels, d.canny, d.gt)
        assert (got[j][0] - out["mask_out"].cpu()).abs().max() <= 1e-3
        pi = out["per_image"].cpu()
        assert (got[j][1] - pi).abs().max() <= 1e-6


def test_host_pipeline_raises_on_graph_overflow(models):
    """a label >= label_cap is dropped by the build; the status word now comes back with the results and
    HostPipeline.synchronize() raises instead of returning metrics of a truncated graph."""
    from camouflage_multimodal_b200 import _lib
    from camouflage_multimodal_b200.pipeline import EvalPipeline, HostPipeline
    gnn, det, kg, _, _ = models
    B = 2
    dev = torch.device("cuda")
    b = synth.make_batch(3, B, seed=77, n_regions=[200, 210])
    ok = EvalPipeline(gnn, det, kg, label_cap=256)
    hp = HostPipeline(ok, B, 256, 256, dev, slots=1, use_graph=True)
    st = hp.submit(0, *(t.pin_memory() for t in (b.images, b.labels, b.canny, b.gt)))
    hp.synchronize()                                                  # fits: no exception
    assert int(st.h_status.item()) == 0
    small = EvalPipeline(gnn, det, kg, label_cap=128)                 # labels 128..210 do not fit
    hp2 = HostPipeline(small, B, 256, 256, dev, slots=1, use_graph=True)
    hp2.submit(0, *(t.pin_memory() for t in (b.images, b.labels, b.canny, b.gt)))
    with pytest.raises(_lib.CmError, match="label >= label_cap"):
        hp2.synchronize()
    d = b.to("cuda")
    out = small.run(d.images, d.labels, d.canny, d.gt)
    assert int(out["status"].item()) & _lib.STATUS_LABEL_OVERFLOW


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_second_device_with_current_device_left_at_zero(models):
    """per-device kernel attributes + a device guard around every library call -- a pipeline on cuda:1 while
    the current device stays 0 gives what cuda:0 gives."""
    from camouflage_multimodal_b200.fusion_model import build_multimodal_model
    from camouflage_multimodal_b200.pipeline import EvalPipeline
    from camouflage_multimodal_b200.region_graph import RegionGraphGNN
    gnn, det, kg, _, _ = models
    b = synth.make_batch(11, 2, seed=5)
    d0 = b.to("cuda:0")
    out0 = EvalPipeline(gnn, det, kg, label_cap=544, lane=11).run(d0.images, d0.labels, d0.canny, d0.gt)
    ref = out0["mask_out"].cpu(), out0["per_image"].cpu()
    torch.cuda.set_device(0)
    gnn1, det1 = RegionGraphGNN(15, 128, 2), build_multimodal_model({})
    gnn1.load_state_dict(gnn.state_dict())
    det1.load_state_dict(det.state_dict())
    gnn1, det1 = gnn1.to("cuda:1").eval(), det1.to("cuda:1").eval()
    d1 = b.to("cuda:1")
    out1 = EvalPipeline(gnn1, det1, kg.to("cuda:1"), label_cap=544, lane=12).run(d1.images, d1.labels, d1.canny, d1.gt)
    assert torch.cuda.current_device() == 0
    torch.cuda.synchronize(1)
    assert torch.equal(out1["mask_out"].cpu(), ref[0]) and torch.equal(out1["per_image"].cpu(), ref[1])


def test_bit_packed_masks_equal_byte_masks(models):
    """Canny + ground truth sent one bit per pixel (BatchBuffer(packed_masks=True) -> cm_unpack_bits inside the captured
    step) give bit-identical results to the byte form."""
    from camouflage_multimodal_b200.pipeline import EvalPipeline, HostPipeline, pack_mask_bits
    gnn, det, kg, _, _ = models
    B = 4
    dev = torch.device("cuda")
    b = synth.make_batch(21, B, seed=1234).as_decoded()
    d = b.to("cuda")
    ref = EvalPipeline(gnn, det, kg, label_cap=544, lane=20).run(d.images, d.labels, d.canny, d.gt)
    ref_mask, ref_per, ref_sod = ref["mask_out"].cpu(), ref["per_image"].cpu(), ref["per_image_sod"].cpu()
    packed = EvalPipeline(gnn, det, kg, label_cap=544, lane=21).run(d.images, d.labels, pack_mask_bits(b.canny).cuda(), pack_mask_bits(b.gt).cuda())
    assert torch.equal(packed["mask_out"].cpu(), ref_mask) and torch.equal(packed["per_image"].cpu(), ref_per)
    hp = HostPipeline(EvalPipeline(gnn, det, kg, label_cap=544, lane=22), B, 256, 256, dev, slots=2, use_graph=True,
                      image_dtype=torch.uint8, label_dtype=torch.int16, packed_masks=True)
    hb = hp.host_batch().fill(b.images, b.labels, b.canny, b.gt)
    assert hb.payload_bytes == B * 65536 * (3 + 2) + 2 * B * 8192
    st = hp.submit(0, hb)
    hp.synchronize()
    assert torch.equal(st.h_mask_out, ref_mask) and torch.equal(st.h_per_image, ref_per)
    assert torch.equal(st.h_per_image_sod[:, :7], ref_sod[:, :7])


def test_image_only_pipeline_equals_explicit_front_end(models):
    """configs[0] "as shipped" in batch form: only uint8 images + ground truth go in, cm_slic / cm_canny run on the device
    inside the (captured) step; identical to running the front end by hand and feeding its maps to the pipeline."""
    from camouflage_multimodal_b200.pipeline import EvalPipeline, HostPipeline
    from camouflage_multimodal_b200.region_graph import canny_edges, slic_segments
    gnn, det, kg, _, _ = models
    B = 4
    dev = torch.device("cuda")
    b = synth.make_batch(31, B, seed=1234).as_decoded()
    d = b.to("cuda")
    lab = slic_segments(d.images)
    can = canny_edges(d.images)
    ref = EvalPipeline(gnn, det, kg, label_cap=544, lane=30).run(d.images, lab, can, d.gt)
    ref_mask, ref_per = ref["mask_out"].cpu(), ref["per_image"].cpu()
    got = EvalPipeline(gnn, det, kg, label_cap=544, lane=31).run(d.images, None, None, d.gt)
    assert torch.equal(got["mask_out"].cpu(), ref_mask) and torch.equal(got["per_image"].cpu(), ref_per)
    assert int(got["status"].item()) == 0
    hp = HostPipeline(EvalPipeline(gnn, det, kg, label_cap=544, lane=32), B, 256, 256, dev, slots=2, use_graph=True,
                      image_dtype=torch.uint8, packed_masks=True, image_only=True)
    hb = hp.host_batch().fill(b.images, None, None, b.gt)
    assert hb.payload_bytes == B * 65536 * 3 + B * 8192 and hb.labels is None
    for j in range(3):                     # slot reuse through the captured graphs
        st = hp.submit(j, hb)
    hp.synchronize()
    assert torch.equal(st.h_mask_out, ref_mask) and torch.equal(st.h_per_image, ref_per)


def test_ten_bit_label_stream_round_trip_and_pipeline():
    """Label maps at 10 bits per pixel (pipeline.pack_labels10 -> cm_unpack_labels10): exact round trip incl. 0 and 1023, and a
    pipeline run on the packed stream gives the outputs of the int16 run bit for bit."""
    import bench
    from camouflage_multimodal_b200 import _lib, synth
    from camouflage_multimodal_b200.pipeline import EvalPipeline, pack_labels10, pack_mask_bits
    dev = torch.device("cuda")
    lib = _lib.load()
    rng = np.random.default_rng(4)
    lab = rng.integers(0, 1024, size=(3, 48, 80)).astype(np.int16)
    lab[0, 0, :4] = [0, 1023, 512, 1]
    packed = pack_labels10(torch.as_tensor(lab))
    assert packed.shape == (3, 48 * 80 * 10 // 8) and packed.dtype == torch.uint8
    out = torch.empty((3, 48, 80), dtype=torch.int16, device=dev)
    _lib.check(lib.cm_unpack_labels10(_lib.ptr(packed.to(dev)), lab.size, _lib.ptr(out), _lib.stream_ptr(dev)), "cm_unpack_labels10")
    assert np.array_equal(out.cpu().numpy(), lab)
    with pytest.raises(ValueError):
        pack_labels10(torch.full((1, 4, 4), 1024))
    gnn, det = bench.make_models()
    kg, _ = bench.load_kg()
    pipe = EvalPipeline(gnn.to(dev), det.to(dev), kg.to(dev), label_cap=bench.LABEL_CAP)
    b = synth.make_batch(0, 6, seed=77, device=dev).as_decoded()
    keys = ("mask_out", "per_image", "per_image_sod", "counts")
    ref = {k: v.clone() for k, v in pipe.run(b.images, b.labels, b.canny, b.gt).items() if k in keys}
    got = pipe.run(b.images, pack_labels10(b.labels).to(dev), pack_mask_bits(b.canny).to(dev), pack_mask_bits(b.gt).to(dev))
    torch.cuda.synchronize()
    for k in keys:
        assert torch.equal(got[k], ref[k]) or torch.allclose(got[k], ref[k], rtol=0, atol=0, equal_nan=True), k


def test_pipeline_is_deterministic_run_to_run_and_across_concurrent_lanes():
    """No floating-point atomics, fixed-order partial sums: the same batch gives bit-identical logits / metrics every time,
    also with two pipelines replaying their captured steps concurrently on separate streams (a missing barrier in the fused
    kernels would show up here as run-to-run noise)."""
    import bench
    from camouflage_multimodal_b200 import synth
    from camouflage_multimodal_b200.pipeline import EvalPipeline
    dev = torch.device("cuda")
    gnn, det = bench.make_models()
    kg, _ = bench.load_kg()
    gnn, det, kg = gnn.to(dev), det.to(dev), kg.to(dev)
    b = synth.make_batch(3, 32, seed=2024, device=dev).as_decoded()
    keys = ("mask_out", "fused", "per_image", "per_image_sod", "counts", "score_out")
    pipes = [EvalPipeline(gnn, det, kg, label_cap=bench.LABEL_CAP, lane=i) for i in range(2)]
    ref = {k: v.clone() for k, v in pipes[0].run(b.images, b.labels, b.canny, b.gt).items() if k in keys}
    for _ in range(4):
        out = pipes[0].run(b.images, b.labels, b.canny, b.gt)
        torch.cuda.synchronize()
        for k in keys:
            assert torch.equal(out[k].view(torch.uint8), ref[k].view(torch.uint8)), k
    graphs = [p.capture(b.images, b.labels, b.canny, b.gt) for p in pipes]
    streams = [torch.cuda.Stream(dev) for _ in pipes]
    torch.cuda.synchronize()
    for _ in range(6):
        for (g, _), s in zip(graphs, streams):
            with torch.cuda.stream(s):
                g.replay()
    torch.cuda.synchronize()
    for _, out in graphs:
        for k in keys:
            assert torch.equal(out[k].view(torch.uint8), ref[k].view(torch.uint8)), k
