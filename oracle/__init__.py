"""CPU oracle for the region-graph -> GNN -> RGxKG fusion -> paint/metrics hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``camouflage_multimodal_b200/`` imports this package;
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may, and there only as the checker or the timed CPU baseline.

Every function restates, in numpy / scipy / plain torch on the CPU, what the reference
(rajan-dubey8/camouflage-multimodal, mounted at /root/reference in the build container) computes on
this path, and cites the reference file:line it follows.

Parity status (also in DESIGN.md):

* fusion + heads (``oracle.fusion``)  and metrics (``oracle.metrics``): PINNED -- checked against the
  reference's own ``models/multimodal/fusion_model.py`` and ``utils/metrics.py`` executed verbatim in
  the build container; outputs committed as ``tests/golden/*.npz`` by ``oracle/make_golden.py``.
* node features / id compaction / edge weights (``oracle.region_graph``): PINNED for everything the
  reference computes itself -- checked against the reference's ``create_region_graph`` executed
  verbatim with its three scikit-image calls injected (scikit-image is not installed).
* RAG adjacency (``oracle.rag``) and the PyG layers (``oracle.gnn``): PARITY UNPINNED -- the
  arithmetic lives in scikit-image (``graph.RAG``) and torch-geometric (``GATConv``/``GCNConv``/
  ``global_mean_pool``), neither vendored nor installed nor version-pinned by the reference
  (requirements: ``scikit-image>=0.21.0``, ``torch-geometric>=2.3.0``), and the reference has no
  tests.  These modules restate the published algorithms of those libraries.
"""
