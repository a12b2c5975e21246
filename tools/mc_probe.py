#!/usr/bin/env python3
"""Probe NVSwitch multicast (NVLS) support through the driver API: the device attribute, and whether a multicast object
can be created for all visible GPUs with a POSIX-fd shareable handle."""
from cuda.bindings import driver as cu

def ck(r):
    if r[0] != cu.CUresult.CUDA_SUCCESS:
        raise RuntimeError(str(r[0]))
    return r[1] if len(r) == 2 else r[1:]

ck(cu.cuInit(0) + (None,))
n = ck(cu.cuDeviceGetCount())
print("devices", n)
for d in range(n):
    dev = ck(cu.cuDeviceGet(d))
    for name in ("CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED", "CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED",
                 "CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED", "CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED"):
        a = getattr(cu.CUdevice_attribute, name)
        print(d, name, ck(cu.cuDeviceGetAttribute(a, dev)))
try:
    dev = ck(cu.cuDeviceGet(0))
    ctx = ck(cu.cuDevicePrimaryCtxRetain(dev))
    ck(cu.cuCtxSetCurrent(ctx) + (None,))
    prop = cu.CUmulticastObjectProp()
    prop.numDevices = max(n, 1)
    prop.size = 1 << 30
    prop.handleTypes = cu.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR.value
    prop.flags = 0
    gran = ck(cu.cuMulticastGetGranularity(prop, cu.CUmulticastGranularity_flags.CU_MULTICAST_GRANULARITY_RECOMMENDED))
    print("multicast granularity", gran)
    h = ck(cu.cuMulticastCreate(prop))
    print("cuMulticastCreate ok", h)
    for d in range(n):
        r = cu.cuMulticastAddDevice(h, ck(cu.cuDeviceGet(d)))
        print("add device", d, r[0])
except Exception as e:
    print("multicast probe failed:", e)
