"""BASELINE configs[4]: stress sweep -- regions per image 64 -> 4096 and KG vocabulary 1k -> 64k rows -- to show where
the gather-bound GNN and the attention-bound fusion cross over.  One JSON line per point (per-kernel ms from the
library's event hooks, device-timed totals).

  python tools/sweep.py [--images 8] [--nodes 64,256,1024,4096] [--kg 13,1024,4096,16384,65536]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=8)
    ap.add_argument("--nodes", default="64,256,1024,4096")
    ap.add_argument("--kg", default="13,1024,4096,16384,65536")
    args = ap.parse_args()
    from camouflage_multimodal_b200 import _lib, synth
    from camouflage_multimodal_b200.fusion_model import build_multimodal_model
    from camouflage_multimodal_b200.region_graph import RegionGraphGNN, build_region_graphs
    from oracle.fusion import fusion_flops   # the SURVEY 8(d) FLOP formula only (no compute)
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    gnn = RegionGraphGNN(15, 128, 2).to(dev).eval()
    det = build_multimodal_model({}).to(dev).eval()
    B = args.images
    for n_regions in [int(v) for v in args.nodes.split(",")]:
        b = synth.make_batch(0, B, seed=77, device=dev, n_regions=n_regions).as_decoded()
        lab32 = b.labels.to(torch.int32)
        cap = n_regions + 8
        graph = {}

        def build():
            graph["g"] = build_region_graphs(b.images, lab32, b.canny, label_cap=cap, want_edge_index=False, out=graph.get("g"))

        t_build = timed(build)
        g = graph["g"]
        n_nodes, n_edges = g.counts()
        res = {}

        def run_gnn():
            res["r"] = gnn.run(g.x, g.csr_row_ptr, g.csr_col, g.edge_w, None, g.node_ptr[B:], g.node_ptr, B, want=("node_emb", "mask_out"))

        t_gnn = timed(run_gnn)
        emb = res["r"]["node_emb"]
        for k_rows in [int(v) for v in args.kg.split(",")]:
            kg = synth.synthetic_kg(k_rows).to(dev)[None]
            try:
                t_fus = timed(lambda: det.forward_ragged(emb, g.node_ptr, B, kg), reps=2)
            except Exception as e:   # report, do not hide: a point the path cannot run is a finding
                print(json.dumps({"regions": n_regions, "kg_rows": k_rows, "error": str(e)[:200]}))
                continue
            flops = B * fusion_flops(n_nodes / B, k_rows)
            print(json.dumps({"regions_per_image": n_regions, "nodes": n_nodes, "edges": n_edges, "kg_rows": k_rows, "images": B,
                              "graph_build_ms": round(t_build, 4), "gnn_ms": round(t_gnn, 4), "fusion_ms": round(t_fus, 4),
                              "fusion_tflops": round(flops / (t_fus * 1e-3) / 1e12, 3),
                              "attention_share_of_flops": round(8.0 * (n_nodes / B) * k_rows * 256 / fusion_flops(n_nodes / B, k_rows), 3),
                              "images_per_s": round(B / ((t_build + t_gnn + t_fus) * 1e-3), 1)}), flush=True)
            del kg
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
