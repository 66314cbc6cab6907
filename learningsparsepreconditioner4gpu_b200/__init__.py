"""B200-native learned-SPAI solve path (GNN -> G, PCG with M^-1 = G G^T + eps I).

Sub-modules are imported lazily: ``datagen`` and ``neural_cg`` host logic import without a GPU,
everything that computes goes through the C-ABI library ``libb200spai.so`` (see ``_lib``), which is
required -- there is no CPU or eager-PyTorch fallback on the product path.

    neural_cg/   host mirror of the reference's package for this path (workspaces, NodeEdgeProcessing, validate, data,
                 dataset folders)                      device.py   DeviceCSR / DevicePCG handles over the C ABI
    linalg.py    the solver object the reference gets from pymathprim       infer.py    the inference benchmark driver
    parallel.py  one large system row-partitioned over the GPUs of a box    batch.py    many small systems in flight per GPU
    systems.py, datagen.py   synthetic systems of the reference's families   csrc/       CUDA kernels + the C ABI
"""
__version__ = "0.1.0"
