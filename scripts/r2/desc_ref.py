"""Diagnostic: the 128 descriptor bytes cuda-python's cuTensorMapEncodeTiled produces for the probe's parameters."""
import torch
from cuda.bindings import driver as cu
a = torch.arange(34 * 160 + 64, dtype=torch.float32, device="cuda")
E = cu.CUtensorMapDataType, cu.CUtensorMapInterleave, cu.CUtensorMapSwizzle, cu.CUtensorMapL2promotion, cu.CUtensorMapFloatOOBfill
for sw in (0, 3):
    r = cu.cuTensorMapEncodeTiled(E[0].CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, a.data_ptr(), [cu.cuuint64_t(160), cu.cuuint64_t(34)],
                                  [cu.cuuint64_t(640)], [cu.cuuint32_t(32), cu.cuuint32_t(8)], [cu.cuuint32_t(1), cu.cuuint32_t(1)],
                                  E[1].CU_TENSOR_MAP_INTERLEAVE_NONE, E[2](sw), E[3].CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                  E[4].CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)
    print("rc", r[0], "sw", sw, "desc:", " ".join("%016x" % int(w) for w in r[1].opaque))
