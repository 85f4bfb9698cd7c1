"""What does this box support for NVLS?  Device attributes + a single-process multicast bring-up over all
visible GPUs (cuMulticastCreate / AddDevice / BindMem / map) followed by a multimem.ld_reduce sanity check
through torch is NOT attempted here: this only reports what the driver answers.  JSON on stdout."""
import json
import os
import sys

from cuda.bindings import driver as drv


def ck(res, what):
    err = res[0]
    if err != drv.CUresult.CUDA_SUCCESS:
        name = drv.cuGetErrorName(err)[1]
        raise RuntimeError(f"{what}: {name}")
    return res[1:] if len(res) > 2 else (res[1] if len(res) == 2 else None)


out = {}
ck(drv.cuInit(0), "cuInit")
n = ck(drv.cuDeviceGetCount(), "count")
out["devices"] = n
A = drv.CUdevice_attribute
attrs = {}
for d in range(n):
    dev = ck(drv.cuDeviceGet(d), "get")
    attrs[d] = {
        "multicast": ck(drv.cuDeviceGetAttribute(A.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev), "a"),
        "fabric": ck(drv.cuDeviceGetAttribute(A.CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED, dev), "a"),
        "posix_fd": ck(drv.cuDeviceGetAttribute(A.CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev), "a"),
        "vmm": ck(drv.cuDeviceGetAttribute(A.CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev), "a"),
    }
out["attrs"] = attrs
try:
    ctxs = []
    for d in range(n):
        dev = ck(drv.cuDeviceGet(d), "get")
        ctxs.append(ck(drv.cuDevicePrimaryCtxRetain(dev), "ctx"))
    size = 2 << 20
    for ht_name in ("POSIX_FILE_DESCRIPTOR", "NONE", "FABRIC"):
        try:
            prop = drv.CUmulticastObjectProp()
            prop.numDevices = n
            prop.size = size
            prop.handleTypes = getattr(drv.CUmemAllocationHandleType, "CU_MEM_HANDLE_TYPE_" + ht_name)
            prop.flags = 0
            gran = ck(drv.cuMulticastGetGranularity(prop, drv.CUmulticastGranularity_flags.CU_MULTICAST_GRANULARITY_RECOMMENDED), "gran")
            out[f"granularity_{ht_name}"] = int(gran)
            prop.size = max(size, int(gran))
            mc = ck(drv.cuMulticastCreate(prop), "cuMulticastCreate")
            for d in range(n):
                dev = ck(drv.cuDeviceGet(d), "get")
                ck(drv.cuMulticastAddDevice(mc, dev), f"AddDevice {d}")
            out[f"create_{ht_name}"] = "ok"
            if ht_name == "POSIX_FILE_DESCRIPTOR":
                fd = ck(drv.cuMemExportToShareableHandle(mc, prop.handleTypes, 0), "export")
                out["export_fd"] = int(fd)
            ck(drv.cuMemRelease(mc), "release")
        except Exception as e:  # noqa: BLE001
            out[f"create_{ht_name}"] = str(e)
except Exception as e:  # noqa: BLE001
    out["error"] = str(e)
print(json.dumps(out))
try:
    import torch
    import torch.distributed._symmetric_memory as sm
    print(json.dumps({"torch_symm_backend": str(sm.get_backend(torch.device('cuda:0'))) if hasattr(sm, 'get_backend') else None}))
except Exception as e:  # noqa: BLE001
    print(json.dumps({"torch_symm_err": str(e)}))
os.system("nvidia-smi topo -m 2>&1 | head -14; nvidia-smi -q 2>/dev/null | grep -i -A3 fabric | head -20; ls /dev/nvidia-caps* /dev/nvidia-nvswitch* 2>&1 | head; uname -r; cat /proc/sys/kernel/yama/ptrace_scope 2>/dev/null")
