#!/usr/bin/env python3
"""Build oracle/_ref/libraycast_ref.so from the reference's OWN sources (test infrastructure).

The reference (a Visual Studio / SDL2 / clew application) cannot be built as a whole here:
states/MainState.cpp needs SDL.h, SDL_ttf.h, clew.h and relies on MSVC binding temporaries to
`Cube(glm::vec4&)`.  What does compile with g++ is exactly the hot path, so this recipe

  * compiles RayTrace/Cube.cpp and RayTrace/lodepng.cpp verbatim, where they lie, with the
    vendored glm (-I/root/reference/RayTrace);
  * extracts, BY FUNCTION NAME, the bodies of MainState::intersectTri / intersectSphere /
    collide / createScene1-3, the ray loop of MainState::executeRayTracerCPU, the intersection
    macros, Utility::normaliseFloat(float) / convertAngleToRadian and Random::getFloat from the
    reference files into a scratch directory under /tmp (never into this repository), and
    compiles them unmodified against the stub declarations in oracle/ref_shim.cpp.

Only the resulting shared object lands in oracle/_ref/ (git-ignored; it travels to the GPU box
with the snapshot).  No reference source text is written anywhere under /root/repo.
Flags: -O2 -ffp-contract=off (no FMA contraction, no fast-math), as DESIGN.md states.
"""
import os
import re
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("RTB_REFERENCE_ROOT", "/root/reference")
RT = os.path.join(REF, "RayTrace")
OUT_DIR = os.path.join(HERE, "_ref")
OUT_SO = os.path.join(OUT_DIR, "libraycast_ref.so")


def _match_braces(text, open_idx):
    depth = 0
    for i in range(open_idx, len(text)):
        c = text[i]
        if c == "{":
            depth += 1
        elif c == "}":
            depth -= 1
            if depth == 0:
                return i
    raise RuntimeError("unbalanced braces")


def extract_function(text, signature_regex):
    """Return the full definition (signature + body) of the first function matching the regex."""
    m = re.search(signature_regex, text)
    if not m:
        raise RuntimeError("signature not found: " + signature_regex)
    open_idx = text.index("{", m.end() - 1)
    close_idx = _match_braces(text, open_idx)
    return text[m.start():close_idx + 1]


def extract_block_after(text, anchor_regex):
    """Return the brace block that starts right after the anchor (anchor text included)."""
    m = re.search(anchor_regex, text)
    if not m:
        raise RuntimeError("anchor not found: " + anchor_regex)
    open_idx = text.index("{", m.end())
    close_idx = _match_braces(text, open_idx)
    return text[m.start():close_idx + 1]


def available():
    return os.path.isfile(os.path.join(RT, "states", "MainState.cpp"))


def build(force=False, quiet=True):
    if not available():
        return None
    srcs = [os.path.join(RT, "states", "MainState.cpp"), os.path.join(RT, "Cube.cpp"),
            os.path.join(RT, "misc", "Utility.cpp"), os.path.join(RT, "misc", "Random.cpp"),
            os.path.join(HERE, "ref_shim.cpp"), os.path.abspath(__file__)]
    if (not force and os.path.isfile(OUT_SO)
            and all(os.path.getmtime(OUT_SO) >= os.path.getmtime(s) for s in srcs)):
        return OUT_SO
    ms = open(srcs[0], encoding="utf-8", errors="replace").read()
    ut = open(srcs[2], encoding="utf-8", errors="replace").read()
    uh = open(os.path.join(RT, "misc", "Utility.h"), encoding="utf-8", errors="replace").read()
    rnd = open(srcs[3], encoding="utf-8", errors="replace").read()

    parts = []
    # intersection macros (EPSILON, CROSS, DOT, SUB): from '#define EPSILON' to the end of SUB
    m0 = ms.index("#define EPSILON")
    m1 = ms.index("#define SUB")
    m1 = ms.index("\n\n", m1)
    parts.append(ms[m0:m1] + "\n")
    pi_line = re.search(r"const float PI\s*=\s*[^;]+;", uh).group(0)
    parts.append("namespace Utility { %s }\n" % pi_line)
    parts.append(extract_function(ut, r"float\s+Utility::normaliseFloat\s*\(\s*float[^)]*\)\s*\{"))
    parts.append(extract_function(ut, r"float\s+Utility::convertAngleToRadian\s*\([^)]*\)\s*\{"))
    parts.append(extract_function(rnd, r"float\s+Random::getFloat\s*\([^)]*\)\s*\{"))
    parts.append(extract_function(ms, r"int\s+MainState::intersectTri\s*\([^)]*\)\s*\{"))
    parts.append(extract_function(ms, r"float\s+MainState::intersectSphere\s*\([^)]*\)\s*\{"))
    parts.append(extract_function(ms, r"glm::vec4\s+MainState::collide\s*\([^)]*\)\s*\{"))
    for n in (1, 2, 3):
        parts.append(extract_function(ms, r"void\s+MainState::createScene%d\s*\(\s*\)\s*\{" % n))
    cpu = extract_function(ms, r"void\s+MainState::executeRayTracerCPU\s*\(\s*\)\s*\{")
    loop = extract_block_after(cpu, r"for\s*\(\s*auto\s+rayOrigin\s*:\s*rayOrigins\s*\)")
    parts.append("void MainState::referenceRayLoop()\n{\n" + loop + "\n}\n")

    os.makedirs(OUT_DIR, exist_ok=True)
    with tempfile.TemporaryDirectory(prefix="rtb_ref_build_") as tmp:
        inc = os.path.join(tmp, "reference_extract.inc")
        with open(inc, "w") as f:
            f.write("\n\n".join(parts))
        common = ["-std=c++14", "-O2", "-ffp-contract=off", "-fPIC", "-w", "-I", RT]
        objs = []
        jobs = [
            (os.path.join(RT, "Cube.cpp"), ["-DCube=RefCube"]),
            (os.path.join(RT, "lodepng.cpp"), []),
            (os.path.join(HERE, "ref_shim.cpp"), ["-I", tmp]),
        ]
        for src, extra in jobs:
            obj = os.path.join(tmp, os.path.basename(src) + ".o")
            cmd = ["g++"] + common + extra + ["-c", src, "-o", obj]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError("reference build failed: %s\n%s" % (" ".join(cmd), r.stderr))
            objs.append(obj)
        r = subprocess.run(["g++", "-shared", "-o", OUT_SO] + objs, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("reference link failed:\n" + r.stderr)
    if not quiet:
        print("built", OUT_SO)
    return OUT_SO


OUT_CL_SO = os.path.join(OUT_DIR, "libraytracer_cl_ref.so")
CL_SRC = os.path.join(RT, "resources", "shaders", "rayTracer.cl")


def build_cl(force=False, quiet=True):
    """oracle/_ref/libraytracer_cl_ref.so: the reference's OpenCL kernel source compiled as C++ for
    the host cores (cl_shim.h + cl_ref_driver.cpp).  The kernel text is read where it lies; the scratch
    copy differs only in the vector-literal syntax `(float4)(...)` -> `make_float4(...)`."""
    if not os.path.isfile(CL_SRC):
        return None
    deps = [CL_SRC, os.path.join(HERE, "cl_shim.h"), os.path.join(HERE, "cl_ref_driver.cpp"), os.path.abspath(__file__)]
    if (not force and os.path.isfile(OUT_CL_SO)
            and all(os.path.getmtime(OUT_CL_SO) >= os.path.getmtime(d) for d in deps)):
        return OUT_CL_SO
    text = open(CL_SRC, encoding="utf-8", errors="replace").read()
    n_lit = len(re.findall(r"\(float4\)\s*\(", text))
    text = re.sub(r"\(float4\)\s*\(", "make_float4(", text)
    os.makedirs(OUT_DIR, exist_ok=True)
    with tempfile.TemporaryDirectory(prefix="rtb_ref_cl_build_") as tmp:
        with open(os.path.join(tmp, "raytracer_cl.inc"), "w") as f:
            f.write(text)
        cmd = ["g++", "-std=c++14", "-O2", "-ffp-contract=off", "-fopenmp", "-fPIC", "-w", "-shared", "-I", HERE, "-I", tmp,
               os.path.join(HERE, "cl_ref_driver.cpp"), "-o", OUT_CL_SO]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("reference OpenCL-kernel build failed: %s\n%s" % (" ".join(cmd), r.stderr))
    if not quiet:
        print("built", OUT_CL_SO, "(%d vector literals rewritten)" % n_lit)
    return OUT_CL_SO


if __name__ == "__main__":
    build_cl(force="--force" in sys.argv, quiet=False)
    so = build(force="--force" in sys.argv, quiet=False)
    if so is None:
        print("reference sources not present at %s; nothing built" % REF)
