"""B200-native learned-SPAI solve path (GNN -> G, PCG with M^-1 = G G^T + eps I).

Sub-modules are imported lazily: ``datagen`` and ``neural_cg`` host logic import without a GPU,
everything that computes goes through the C-ABI library ``libb200spai.so`` (see ``_lib``), which is
required -- there is no CPU or eager-PyTorch fallback on the product path.
"""
__version__ = "0.1.0"
