"""Builds and runs the C++ host-side mirror's tests (tests/cpp_mirror_test.cpp over include/image_b200.hpp)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "image_b200", "lib", "cpp_mirror_test")


def _build():
    lib = os.path.join(ROOT, "image_b200", "lib")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cpp_mirror_test.cpp"), "-o", EXE, "-L", lib, "-lb200img",
                           f"-Wl,-rpath,{lib}"])


def test_cpp_mirror_compiles_and_links():
    _build()
    assert os.path.exists(EXE)


@pytest.mark.gpu
def test_cpp_mirror_runs(cuda):
    _build()
    r = subprocess.run([EXE], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "cpp mirror ok" in r.stdout
