from cuda.bindings import driver as drv
import ctypes
drv.cuInit(0)
err, dev = drv.cuDeviceGet(0)
err, ctx = drv.cuDevicePrimaryCtxRetain(dev)
drv.cuCtxSetCurrent(ctx)
err, d = drv.cuMemAlloc(752*480*3)
print("alloc", err, hex(int(d)))
r = drv.cuTensorMapEncodeTiled(drv.CUtensorMapDataType.CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, int(d) + 752*480,
        [drv.cuuint64_t(752), drv.cuuint64_t(480)], [drv.cuuint64_t(752)], [drv.cuuint32_t(64), drv.cuuint32_t(64)], [drv.cuuint32_t(1), drv.cuuint32_t(1)],
        drv.CUtensorMapInterleave.CU_TENSOR_MAP_INTERLEAVE_NONE, drv.CUtensorMapSwizzle.CU_TENSOR_MAP_SWIZZLE_NONE,
        drv.CUtensorMapL2promotion.CU_TENSOR_MAP_L2_PROMOTION_NONE, drv.CUtensorMapFloatOOBfill.CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)
print(r[0])
tm = r[1]
print("py desc", " ".join("%016x" % int(x) for x in tm.opaque))
