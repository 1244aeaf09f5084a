"""CPU (-m "not gpu"): the header-only C++ drop-in (shim/*.h) and every native driver under tests/native compile with
g++ and link against libcuorb.so without a GPU - a maintainer's build of the reference with the shim swapped in must
not depend on the box it is built on.  (What the drivers compute is checked on the GPU by tests/test_gpu_shim.py.)"""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "orbslam3_cpp_fork_b200")
PLAIN = ["shim_check.cpp", "prep_check.cpp", "search_check.cpp"]                      # own minimal Frame / KeyFrame structs
OVER_REF_NAMES = ["lastframe_check.cpp", "rig_check.cpp", "fisheye_check.cpp",        # the stand-in classes of oracle/ref_compat
                  "localmap_check.cpp", "localmap_resident_check.cpp"]                # (the reference's member names)


@pytest.fixture(scope="module")
def lib():
    from orbslam3_cpp_fork_b200 import build
    build.build_libcuorb()
    return PKG


@pytest.mark.parametrize("src", PLAIN + OVER_REF_NAMES)
def test_native_driver_compiles_and_links(lib, tmp_path, src):
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Werror=return-type"]
    if src in OVER_REF_NAMES:
        cmd += ["-ffp-contract=off", "-DCUORB_SHIM_USE_OPENCV", "-I" + os.path.join(ROOT, "oracle", "ref_compat")]
    exe = str(tmp_path / src.replace(".cpp", ""))
    cmd += [os.path.join(ROOT, "tests", "native", src), "-o", exe, "-L" + lib, "-l:libcuorb.so", "-Wl,-rpath," + lib]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    # no arguments: every driver prints its usage and exits non-zero before touching the device
    u = subprocess.run([exe], capture_output=True, text=True)
    assert u.returncode != 0 and "usage" in (u.stderr + u.stdout).lower()


def test_every_shim_header_is_self_contained(tmp_path):
    for h in sorted(os.listdir(os.path.join(ROOT, "shim"))):
        if not h.endswith(".h"):
            continue
        tu = tmp_path / (h + ".cpp")
        tu.write_text('#include "%s"\nint main() { return 0; }\n' % os.path.join(ROOT, "shim", h))
        r = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", str(tu)], capture_output=True, text=True)
        assert r.returncode == 0, h + ":\n" + r.stderr[-3000:]
