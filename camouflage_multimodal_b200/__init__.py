"""B200-native (sm_100a) implementation of the camouflage-multimodal region-graph hot path.

Drop-in surface (same names and signatures as the reference):
  region_graph.create_region_graph / RegionGraphGNN          (models/region_graph/extract_rg_embeddings.py)
  fusion_model.CrossAttentionFusion / LateFusion / MultimodalCamouflageDetector / build_multimodal_model
                                                              (models/multimodal/fusion_model.py)
  metrics.calculate_* / evaluate_segmentation / batch_evaluate (utils/metrics.py)
plus the batched device-resident pipeline (pipeline.EvalPipeline).  Everything computes through
libcm_b200.so (C ABI in include/cm_b200.h); importing this package does not need a GPU, calling it does.
"""
from .data import Batch, Data  # noqa: F401

__all__ = ["Data", "Batch"]
