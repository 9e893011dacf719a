"""The C-ABI library loads on a CPU-only box and exports every symbol include/rtb.h declares; the
host-only entry points behave; device entry points FAIL LOUDLY without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from parity import ROOT


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "rtb.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rtb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_all_exported(built_libs):
    from opencl_ray_tracer_b200 import capi
    lib = capi.load_library()
    declared = _declared_symbols()
    assert len(declared) >= 17
    assert sorted(capi.EXPORTED_SYMBOLS) == declared
    for name in declared:
        assert getattr(lib, name) is not None, name
    assert b"sm_100a" in lib.rtb_version()


def test_library_is_sm100a_only(built_libs):
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", os.path.join(built_libs, "librtb.so")], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    assert "sm_100a" in out.stdout


def test_error_strings_and_band_rows(built_libs):
    from opencl_ray_tracer_b200 import capi
    lib = capi.load_library()
    for code, name in capi.STATUS_NAMES.items():
        assert lib.rtb_error_string(code).decode() == name
    assert lib.rtb_error_string(99).decode() == "RTB_ERR_UNKNOWN"
    # interleaved bands: 4320 rows, 16-row bands over 8 ranks -> 270 bands, ranks 0..5 own 34, 6..7 own 33
    rows = [lib.rtb_band_local_rows(4320, 16, r, 8) for r in range(8)]
    assert sum(rows) == 4320 and rows[0] == 34 * 16 and rows[7] == 33 * 16
    rows = [lib.rtb_band_local_rows(1083, 16, r, 4) for r in range(4)]  # ragged last band (11 rows)
    assert sum(rows) == 1083
    assert lib.rtb_band_local_rows(0, 16, 0, 1) == 0 and lib.rtb_band_local_rows(10, 16, 2, 2) == 0


def test_no_cpu_fallback_without_device(built_libs):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from opencl_ray_tracer_b200 import capi
    with pytest.raises(capi.RtbError) as ei:
        capi.Context(0)
    assert ei.value.status == 2  # RTB_ERR_NO_DEVICE


def test_missing_library_fails_loudly(tmp_path):
    from opencl_ray_tracer_b200 import capi
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        capi.load_library(str(tmp_path / "librtb.so"))


def test_packer_rejects_non_boxes_and_bad_cameras(built_libs):
    from opencl_ray_tracer_b200 import capi, host_scene as hs
    s = hs.reference_scene(1)
    cr, sr, cam = capi.debug_pack(s.sph_o, s.sph_r, s.cube_v)
    assert cr.shape == (4, 3, 4) and sr.shape == (2, 4) and np.isfinite(cr).all()
    assert cam[0] == 1.0 and cam[1] == 0.0 and cam[2] == 1.0  # u = x, v = y for the reference camera
    bad = s.cube_v.copy()
    bad[1, 7, 0] += 3.0
    with pytest.raises(capi.RtbError) as ei:
        capi.debug_pack(s.sph_o, s.sph_r, bad)
    assert ei.value.status == 5 and "cube 1" in str(ei.value)
    o0, du, dv, d = capi.REFERENCE_CAMERA
    with pytest.raises(capi.RtbError) as ei:
        capi.debug_pack(s.sph_o, s.sph_r, s.cube_v, (o0, du, du, d))  # dU parallel to dV
    assert ei.value.status == 6


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "opencl_ray_tracer_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".h", ".cuh", ".sh")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                # comments may MENTION the oracle; nothing may import, include, link or open it
                hits = re.findall(r"(?:import\s+oracle|from\s+oracle|#include[^\n]*oracle|libraycast_(?:oracle|ref)"
                                  r"|-loracle|oracle/_ref/|oracle\.py)", text)
                hits = [h for h in hits if not (f.endswith((".cpp", ".h")) and h == "oracle/_ref/")]
                assert not hits, (os.path.join(dirpath, f), hits)


def test_header_is_plain_c(tmp_path):
    """include/rtb.h is the boundary a C or C++ host binds: it must compile as C99 and as C++11 on its own."""
    import subprocess
    src = tmp_path / "hdr.c"
    src.write_text('#include "rtb.h"\nint main(void) { rtb_ctx* c = 0; (void)c; return RTB_OK; }\n')
    inc = os.path.join(ROOT, "include")
    subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-I", inc, "-fsyntax-only", str(src)], check=True)
    subprocess.run(["g++", "-std=c++11", "-Wall", "-Werror", "-I", inc, "-fsyntax-only", "-x", "c++", str(src)], check=True)
