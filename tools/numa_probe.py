"""Host topology facts behind the multi-GPU e2e numbers: NUMA nodes, the node of every GPU's PCIe slot, CPU affinity of
this process, and whether this container may call mbind / set_mempolicy (page placement of the shared host frame)."""
import ctypes
import glob
import os
import subprocess

print("nodes:", sorted(os.path.basename(p) for p in glob.glob("/sys/devices/system/node/node[0-9]*")))
for p in sorted(glob.glob("/sys/devices/system/node/node[0-9]*/cpulist")):
    print(p, open(p).read().strip())
for p in sorted(glob.glob("/sys/devices/system/node/node[0-9]*/meminfo")):
    print(open(p).read().splitlines()[0])
print("affinity:", sorted(os.sched_getaffinity(0)))
print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout)
out = subprocess.run(["nvidia-smi", "--query-gpu=index,pci.bus_id", "--format=csv,noheader"], capture_output=True, text=True).stdout
for ln in out.strip().splitlines():
    idx, bdf = [x.strip() for x in ln.split(",")]
    f = "/sys/bus/pci/devices/%s/numa_node" % bdf.lower().replace("00000000:", "0000:")
    print("gpu", idx, bdf, "numa_node", open(f).read().strip() if os.path.exists(f) else "?")
libc = ctypes.CDLL(None, use_errno=True)
SYS_mbind, SYS_get_mempolicy = 237, 239
buf = ctypes.create_string_buffer(1 << 20)
addr = (ctypes.addressof(buf) + 4095) & ~4095
mask = ctypes.c_ulong(1)
rc = libc.syscall(SYS_mbind, ctypes.c_void_p(addr), ctypes.c_ulong(1 << 16), 2, ctypes.byref(mask), ctypes.c_ulong(64), 0)  # MPOL_BIND node 0
print("mbind(MPOL_BIND, node 0) rc", rc, "errno", ctypes.get_errno(), os.strerror(ctypes.get_errno()) if rc else "")
mask = ctypes.c_ulong(3)
rc = libc.syscall(SYS_mbind, ctypes.c_void_p(addr), ctypes.c_ulong(1 << 16), 3, ctypes.byref(mask), ctypes.c_ulong(64), 0)  # MPOL_INTERLEAVE nodes 0,1
print("mbind(MPOL_INTERLEAVE, nodes 0-1) rc", rc, "errno", ctypes.get_errno(), os.strerror(ctypes.get_errno()) if rc else "")
