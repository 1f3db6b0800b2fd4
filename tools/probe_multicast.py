"""Prints what the box offers for NVSwitch multicast (Path B): run on the GPU box."""
from cuda.bindings import driver as d


def main():
    (err,) = d.cuInit(0)
    err, n = d.cuDeviceGetCount()
    A = d.CUdevice_attribute
    for i in range(n):
        err, dev = d.cuDeviceGet(i)
        vals = {}
        for name in ("CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED", "CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED",
                     "CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED",
                     "CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED"):
            err, v = d.cuDeviceGetAttribute(getattr(A, name), dev)
            vals[name.replace("CU_DEVICE_ATTRIBUTE_", "")] = v
        print("device", i, vals)
    if n >= 2:
        prop = d.CUmulticastObjectProp()
        prop.numDevices = n
        prop.size = 2 << 20
        prop.handleTypes = d.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR
        err, g1 = d.cuMulticastGetGranularity(prop, d.CUmulticastGranularity_flags.CU_MULTICAST_GRANULARITY_MINIMUM)
        err2, g2 = d.cuMulticastGetGranularity(prop, d.CUmulticastGranularity_flags.CU_MULTICAST_GRANULARITY_RECOMMENDED)
        print("multicast granularity min/recommended:", err, g1, err2, g2)


if __name__ == "__main__":
    main()
