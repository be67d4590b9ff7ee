#!/usr/bin/env python
"""Prints what the box says about how its GPUs reach the host: NVML common-ancestor levels, PCI bus
ids, sysfs device paths, NUMA/CPU affinity, and the placement `launch.spread_devices` derives."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import pynvml
    import torch
    import b200dp  # noqa: F401
    from b200dp.launch import pcie_closeness, spread_devices
    pynvml.nvmlInit()
    n = torch.cuda.device_count()
    print("visible GPUs:", n, "| CUDA_VISIBLE_DEVICES =", os.environ.get("CUDA_VISIBLE_DEVICES"))
    hs = []
    for i in range(n):
        h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(torch.cuda.get_device_properties(i).uuid)).encode())
        hs.append(h)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        name = bus.lower()[-12:]
        path = os.path.realpath(f"/sys/bus/pci/devices/{name}")
        numa = "?"
        try:
            numa = open(f"/sys/bus/pci/devices/{name}/numa_node").read().strip()
        except OSError:
            pass
        try:
            gen = pynvml.nvmlDeviceGetCurrPcieLinkGeneration(h)
            width = pynvml.nvmlDeviceGetCurrPcieLinkWidth(h)
        except Exception:
            gen = width = "?"
        print(f"gpu{i}: bus {bus} numa {numa} pcie gen{gen} x{width} sysfs {path}")
    print("NVML common ancestor level (0 internal, 10 single switch, 20 multiple switches, 30 host bridge, 40 node, 50 system):")
    for i in range(n):
        row = []
        for j in range(n):
            if i == j:
                row.append("  -")
                continue
            try:
                row.append(f"{int(pynvml.nvmlDeviceGetTopologyCommonAncestor(hs[i], hs[j])):3d}")
            except Exception as e:
                row.append("  ?")
        print("  ", " ".join(row))
    c = pcie_closeness()
    print("closeness matrix (NVML / sysfs):", c)
    if c is None and n > 2:
        from b200dp.launch import measured_spread
        for w in (2, 4):
            print(f"measured placement for {w} ranks:", measured_spread(w))
        import glob, json, tempfile
        for f in sorted(glob.glob(os.path.join(tempfile.gettempdir(), "b200dp_pcie_*.json"))):
            print(f, open(f).read()[:1500])
    for w in (2, 4):
        print(f"spread placement for {w} ranks:", spread_devices(w, c, n))


if __name__ == "__main__":
    main()
