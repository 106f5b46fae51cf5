"""The reference's CPU path for the whole hot path, one image at a time (oracle; test infrastructure only --
timed by bench.py's cpu_baseline / --impl reference legs, never shipped).

Per image, exactly what the reference does across extract_embeddings_from_image
(models/region_graph/extract_rg_embeddings.py:249-288), predict_single_image
(models/multimodal/test_multimodal.py:83-152) and detect_camouflage + utils/metrics
(models/region_graph/test.py:236-293): per-region Python loop over full-frame masks (``faithful=True``, the
reference's own algorithm and cost), RAG, GNN forward, B=1 fusion forward, paint, metrics.
"""
from __future__ import annotations

import time

import numpy as np
import torch

from . import fusion as ofus
from . import gnn as ognn
from . import metrics as om
from . import region_graph as org


def run_image(image, labels, canny, gt_u8, gnn_sd, fusion_sd, kg, faithful=True, sod=False):
    """-> (mask logits [2], metrics dict) for one image; ``kg`` is ``[K,128]``.  ``sod``: also S- / E- / F- / weighted
    F-measure of the painted mask (oracle/sod_metrics.py; not in the reference)."""
    x, ei, w, _ = org.build_region_graph(np.asarray(image, np.float64), np.asarray(labels, np.int64),
                                         np.asarray(canny).astype(bool), faithful=faithful)
    with torch.no_grad():
        r = ognn.rg_gnn_forward(gnn_sd, torch.from_numpy(x), torch.from_numpy(ei), torch.from_numpy(w))
        mask, _, _, _, _, _ = ofus.detector_forward(fusion_sd, r["node_emb"][None], kg[None])
    pred = om.paint_mask(r["mask_out"].numpy(), np.asarray(labels))
    m = om.evaluate_segmentation(pred, np.asarray(gt_u8) / 255.0)
    if sod:
        from . import sod_metrics as osod
        m.update(osod.evaluate_sod(pred, np.asarray(gt_u8) / 255.0)[0])
    return mask[0].numpy(), m


def _worker(args):
    torch.set_num_threads(1)
    return run_image(*args)


def time_images(batch, gnn_sd, fusion_sd, kg, faithful=True, processes=1, sod=False):
    """Seconds to push every image of ``batch`` (a synth.SynthBatch on the CPU) through the path, using
    ``processes`` worker processes (the reference itself is a single-threaded Python loop)."""
    items = [(batch.images[i].numpy(), batch.labels[i].numpy(), batch.canny[i].numpy(), batch.gt[i].numpy(),
              gnn_sd, fusion_sd, kg, faithful, sod) for i in range(len(batch))]
    t0 = time.perf_counter()
    if processes <= 1:
        out = [run_image(*it) for it in items]
    else:
        import multiprocessing as mp
        with mp.get_context("fork").Pool(processes) as pool:
            out = pool.map(_worker, items, chunksize=1)
    return time.perf_counter() - t0, out
