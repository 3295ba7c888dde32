#!/usr/bin/env python
"""Probe: does this box support NVLink multicast objects (NVLS)?  Prints the
device attribute, the multicast granularity and whether a multicast object
for all visible GPUs can be created and exported as a POSIX fd."""
import sys
from cuda.bindings import driver as cu


def ck(r):
    err = r[0]
    if err != cu.CUresult.CUDA_SUCCESS:
        raise RuntimeError(str(cu.cuGetErrorString(err)[1]))
    return r[1:] if len(r) > 2 else (r[1] if len(r) == 2 else None)


ck(cu.cuInit(0))
n = ck(cu.cuDeviceGetCount())
print("devices", n)
for d in range(n):
    dev = ck(cu.cuDeviceGet(d))
    sup = ck(cu.cuDeviceGetAttribute(
        cu.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev))
    fab = ck(cu.cuDeviceGetAttribute(
        cu.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED, dev))
    print("dev", d, "multicast_supported", sup, "fabric_handles", fab)
dev = ck(cu.cuDeviceGet(0))
ctx = ck(cu.cuDevicePrimaryCtxRetain(dev))
ck(cu.cuCtxSetCurrent(ctx))
prop = cu.CUmulticastObjectProp()
prop.numDevices = n
prop.size = 64 << 20
prop.handleTypes = cu.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR
prop.flags = 0
try:
    gran = ck(cu.cuMulticastGetGranularity(
        prop, cu.CUmulticastGranularity_flags.CU_MULTICAST_GRANULARITY_RECOMMENDED))
    print("granularity", gran)
    h = ck(cu.cuMulticastCreate(prop))
    print("created multicast object", h)
    fd = ck(cu.cuMemExportToShareableHandle(
        h, cu.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0))
    print("exported fd", fd)
except Exception as e:
    print("FAILED:", e)
    sys.exit(0)
