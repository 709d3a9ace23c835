import torch
from cuda.bindings import driver as cu
x = torch.zeros(4, 32, 64, 512, device="cuda")
def enc(box0, es0, sw):
    r = cu.cuTensorMapEncodeTiled(cu.CUtensorMapDataType.CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, x.data_ptr(),
        [cu.cuuint64_t(v) for v in (512, 64, 32, 4)], [cu.cuuint64_t(v) for v in (512*4, 512*64*4, 512*64*32*4)],
        [cu.cuuint32_t(v) for v in (box0, 1, 32, 1)], [cu.cuuint32_t(v) for v in (es0, 1, 1, 1)],
        cu.CUtensorMapInterleave.CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
        cu.CUtensorMapL2promotion.CU_TENSOR_MAP_L2_PROMOTION_L2_256B, cu.CUtensorMapFloatOOBfill.CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)
    return r[0]
SW = cu.CUtensorMapSwizzle
for box0, es0, sw in [(32,1,SW.CU_TENSOR_MAP_SWIZZLE_128B),(64,2,SW.CU_TENSOR_MAP_SWIZZLE_128B),(63,2,SW.CU_TENSOR_MAP_SWIZZLE_128B),(32,2,SW.CU_TENSOR_MAP_SWIZZLE_128B),(64,2,SW.CU_TENSOR_MAP_SWIZZLE_NONE),(64,1,SW.CU_TENSOR_MAP_SWIZZLE_128B)]:
    print(box0, es0, sw, enc(box0, es0, sw))
