"""Does cuTensorMapEncodeTiled accept a dimension whose stride is smaller than the extent of the dimension below it
(overlapping windows)?  Needed for reading 4-pixel windows of a 16-channel image at a 1-pixel step."""
import ctypes, sys
import torch
from cuda.bindings import driver as cu
torch.zeros(1, device="cuda")
buf = torch.zeros(4 * 115 * 115 * 16, dtype=torch.bfloat16, device="cuda")
def enc(rank, dims, strides, box):
    r = cu.cuTensorMapEncodeTiled(cu.CUtensorMapDataType.CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, buf.data_ptr(),
                                  [cu.cuuint64_t(d) for d in dims], [cu.cuuint64_t(s) for s in strides],
                                  [cu.cuuint32_t(b) for b in box], [cu.cuuint32_t(1)] * rank,
                                  cu.CUtensorMapInterleave.CU_TENSOR_MAP_INTERLEAVE_NONE,
                                  cu.CUtensorMapSwizzle.CU_TENSOR_MAP_SWIZZLE_128B,
                                  cu.CUtensorMapL2promotion.CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                  cu.CUtensorMapFloatOOBfill.CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)
    return r[0]
print("A rank4 overlap {64,112,115,N} strides {32,3680,423200}:", enc(4, [64, 112, 115, 4], [32, 3680, 423200], [64, 128, 1, 1]))
print("A' box 112:", enc(4, [64, 112, 115, 4], [32, 3680, 423200], [64, 112, 1, 1]))
print("B rank5 {16,4,112,115,N} strides {32,32,3680,423200}:", enc(5, [16, 4, 112, 115, 4], [32, 32, 3680, 423200], [16, 4, 128, 1, 1]))
