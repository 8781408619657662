"""Builds tests/cpp/host_mirror_test.cpp against include/voda_host.hpp + libvoda_accel.so and
runs it: the C++ mirror of SchedulerAlgorithm / ComputeMunkresMax on the KATs."""
import os
import subprocess

import pytest

from vodascheduler_b200 import build as vbuild

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(vbuild.HERE)


def test_cpp_host_mirror(tmp_path):
    so = vbuild.build()
    exe = str(tmp_path / "host_mirror_test")
    cmd = ["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", "host_mirror_test.cpp"), "-o", exe,
           "-L", os.path.dirname(so), "-lvoda_accel", "-Wl,-rpath," + os.path.dirname(so),
           "-Wl,-rpath,/usr/local/cuda/lib64"]
    subprocess.run(cmd, check=True)
    p = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stdout + p.stderr
    assert "host mirror OK" in p.stdout
