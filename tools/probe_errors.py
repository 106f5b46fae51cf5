"""Prints max-abs errors of the GPU path against the committed golden vectors (run on the GPU box)."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_oracle_golden import fusion_state_dict, gnn_state_dict
from camouflage_multimodal_b200.region_graph import RegionGraphGNN
from camouflage_multimodal_b200.fusion_model import build_multimodal_model
from camouflage_multimodal_b200.data import Data
G = os.path.join(ROOT, "tests", "golden")
z = np.load(os.path.join(G, "region_graph.npz"))
m = RegionGraphGNN(15, 128, 2); m.load_state_dict(gnn_state_dict(z)); m = m.cuda().eval()
for name in "AB":
    d = Data(x=torch.from_numpy(z[f"{name}.x"]), edge_index=torch.from_numpy(z[f"{name}.edge_index"]),
             edge_attr=torch.from_numpy(z[f"{name}.edge_attr"])).to("cuda")
    mask, inst, edge = m(d)
    emb = m.extract_node_embeddings(d)
    ge = m.extract_graph_embedding(d)
    print("gnn", name, "mask", np.abs(mask.cpu().numpy() - z[f"{name}.mask_out"]).max(),
          "inst", np.abs(inst.cpu().numpy() - z[f"{name}.instance_out"]).max(),
          "edge", np.abs(edge.cpu().numpy() - z[f"{name}.edge_out"]).max(),
          "emb", np.abs(emb.cpu().numpy()[::16] - z[f"{name}.node_emb_rows"]).max(), "embmax", float(emb.max()),
          "gemb", np.abs(ge.cpu().numpy() - z[f"{name}.graph_emb"]).max())
z = np.load(os.path.join(G, "fusion.npz"))
det = build_multimodal_model({}); det.load_state_dict(fusion_state_dict(z)); det = det.cuda().eval()
for case in ["single_4d", "batch_3d", "graph_2d", "view_4d"]:
    rg, kg = torch.from_numpy(z[f"{case}.rg"]).cuda(), torch.from_numpy(z[f"{case}.kg"]).cuda()
    mask, inst, edge, score, attn = det(rg, kg, return_attention=True)
    fused, _ = det.fusion(rg, kg)
    print("fusion", case, tuple(rg.shape), tuple(kg.shape), {n: float(np.abs(g.cpu().numpy() - z[f"{case}.{n}"]).max()) for n, g in
          (("mask", mask), ("instance", inst), ("edge", edge), ("score", score), ("fused", fused), ("rg2kg", attn["rg2kg"]), ("kg2rg", attn["kg2rg"]))},
          "fused_absmax", float(np.abs(z[f"{case}.fused"]).max()))
