#!/usr/bin/env python3
"""One-shot probe of the GPU box for a COMPUTE OpenCL ICD: is there a driver library
(libnvidia-opencl.so.1) behind the CUDA toolkit's ICD loader (libOpenCL.so.1), and does clGetPlatformIDs
see a platform / GPU device?  Output is committed as profiles/r02_opencl_icd_probe.txt.  Nothing graphical."""
import ctypes as C
import glob
import os
import subprocess
import sys
import tempfile


def sh(cmd):
    r = subprocess.run(cmd, shell=True, capture_output=True, text=True)
    print("$ %s\n%s%s[rc=%d]" % (cmd, r.stdout, r.stderr, r.returncode))


def try_loader(name, label):
    try:
        lib = C.CDLL(name)
    except OSError as e:
        print("%s: dlopen(%s) failed: %s" % (label, name, e))
        return False
    try:
        n = C.c_uint(0)
        rc = lib.clGetPlatformIDs(0, None, C.byref(n))
    except AttributeError:
        print("%s: %s has no clGetPlatformIDs" % (label, name))
        return False
    print("%s: dlopen(%s) ok; clGetPlatformIDs -> rc=%d, platforms=%d" % (label, name, rc, n.value))
    if rc != 0 or n.value == 0:
        return False
    plats = (C.c_void_p * n.value)()
    lib.clGetPlatformIDs(n.value, plats, None)
    ok = False
    for p in plats:
        buf = C.create_string_buffer(256)
        for what, code in (("name", 0x0902), ("vendor", 0x0903), ("version", 0x0901)):
            lib.clGetPlatformInfo(C.c_void_p(p), code, 256, buf, None)
            print("   platform %s: %s" % (what, buf.value.decode(errors="replace")))
        nd = C.c_uint(0)
        rc = lib.clGetDeviceIDs(C.c_void_p(p), C.c_ulong(4), 0, None, C.byref(nd))   # CL_DEVICE_TYPE_GPU
        print("   clGetDeviceIDs(GPU) -> rc=%d, devices=%d" % (rc, nd.value))
        if rc == 0 and nd.value:
            devs = (C.c_void_p * nd.value)()
            lib.clGetDeviceIDs(C.c_void_p(p), C.c_ulong(4), nd.value, devs, None)
            lib.clGetDeviceInfo(C.c_void_p(devs[0]), 0x102B, 256, buf, None)       # CL_DEVICE_NAME
            print("   device 0: %s" % buf.value.decode(errors="replace"))
            ok = True
    return ok


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "child":
        sys.exit(0 if try_loader(sys.argv[2], sys.argv[3]) else 1)
    sh("ls -la /usr/lib/x86_64-linux-gnu/libnvidia-opencl* /usr/lib64/libnvidia-opencl* /etc/OpenCL/vendors/ 2>&1")
    sh("ls -la /usr/local/cuda/targets/x86_64-linux/lib/libOpenCL* 2>&1")
    sh("ldconfig -p | grep -i -E 'opencl|nvidia-ptxjit|nvidia-nvvm' ")
    sh("which clinfo && clinfo -l")
    sh("nvidia-smi --query-gpu=name,driver_version --format=csv")
    found = False
    loaders = ["libOpenCL.so.1", "/usr/local/cuda/targets/x86_64-linux/lib/libOpenCL.so.1"]
    icds = sorted(set(glob.glob("/usr/lib/x86_64-linux-gnu/libnvidia-opencl.so*") + glob.glob("/usr/lib64/libnvidia-opencl.so*")))
    # every attempt in a child process: a loader that aborts must not take the probe down
    for ld in loaders:
        r = subprocess.run([sys.executable, __file__, "child", ld, "loader, default ICD search"], capture_output=True, text=True)
        print(r.stdout, r.stderr[-500:])
        found = found or r.returncode == 0
        for icd in icds:
            with tempfile.TemporaryDirectory() as d:
                open(os.path.join(d, "nvidia.icd"), "w").write(icd + "\n")
                env = dict(os.environ, OCL_ICD_VENDORS=d, OPENCL_VENDOR_PATH=d, OCL_ICD_FILENAMES=icd)
                r = subprocess.run([sys.executable, __file__, "child", ld, "loader with OCL_ICD_VENDORS -> %s" % icd],
                                   capture_output=True, text=True, env=env)
                print(r.stdout, r.stderr[-500:])
                found = found or r.returncode == 0
    for icd in icds:
        r = subprocess.run([sys.executable, __file__, "child", icd, "driver library opened directly"], capture_output=True, text=True)
        print(r.stdout, r.stderr[-500:])
        found = found or r.returncode == 0
    print("VERDICT: %s" % ("a compute OpenCL platform with a GPU device is usable on this box" if found else
                           "OpenCL ICD unavailable on this box (no driver library / no platform)"))


if __name__ == "__main__":
    main()
