import os
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
if os.path.dirname(__file__) not in sys.path:
    sys.path.insert(0, os.path.dirname(__file__))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def port():
    from oracle.oracle import PortOracle
    return PortOracle()


@pytest.fixture(scope="session")
def ref():
    """The reference's own compiled functions; present only where oracle/_ref was built."""
    from oracle import build_ref
    from oracle.oracle import RefOracle, ref_available
    try:
        build_ref.build()
    except Exception as e:  # pragma: no cover
        pytest.skip("oracle/_ref could not be built: %s" % e)
    if not ref_available():
        pytest.skip("oracle/_ref/libraycast_ref.so not present (reference sources absent)")
    return RefOracle()


@pytest.fixture(scope="session")
def cl_ref():
    """The reference's OpenCL kernel source compiled as C++ for the host; present only where oracle/_ref was built."""
    from oracle import build_ref
    from oracle.oracle import ClRefOracle, cl_ref_available
    try:
        build_ref.build_cl()
    except Exception as e:  # pragma: no cover
        pytest.skip("oracle/_ref (OpenCL kernel) could not be built: %s" % e)
    if not cl_ref_available():
        pytest.skip("oracle/_ref/libraytracer_cl_ref.so not present (reference sources absent)")
    return ClRefOracle()


@pytest.fixture(scope="session")
def built_libs():
    """Makes sure librtb.so / librtb_host.so exist (cross-compiles without a GPU)."""
    import subprocess
    pkg = os.path.join(ROOT, "opencl_ray_tracer_b200")
    if not (os.path.isfile(os.path.join(pkg, "librtb.so")) and os.path.isfile(os.path.join(pkg, "librtb_host.so"))):
        subprocess.run([os.path.join(pkg, "build.sh")], check=True, capture_output=True)
    return pkg


@pytest.fixture(scope="session")
def emu(built_libs):
    from parity import Emulator
    return Emulator()


@pytest.fixture(scope="session")
def gpu_ctx(built_libs):
    from opencl_ray_tracer_b200 import capi
    ctx = capi.Context(0)  # raises loudly without a device: there is no fallback
    yield ctx
    ctx.close()
