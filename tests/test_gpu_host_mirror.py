"""The C++ host mirror used from C++: tests/host_mirror_demo.cpp is compiled against include/dso_b200.h +
host/dso_b200_host.h, linked with libdso_b200.so and run on the GPU (makeImages -> setCoarseTrackingRef ->
trackNewestCoarse -> calcRes / calcGSSSE through the reference's class surface); it must recover the known motion."""
import os
import subprocess
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def test_host_mirror_cpp_program(tmp_path):
    libdir = os.path.join(ROOT, "dso-noted-in-chinese_b200")
    exe = str(tmp_path / "host_mirror_demo")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-I", ROOT, os.path.join(ROOT, "tests", "host_mirror_demo.cpp"),
                           "-o", exe, "-L", libdir, "-ldso_b200", "-Wl,-rpath," + libdir])
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    print(out.stdout, out.stderr)
    assert out.returncode == 0 and "OK" in out.stdout, out.stdout + out.stderr
