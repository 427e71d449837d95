"""Builds tests/reference_suite.cpp (the reference's tests.rs restated against the header-only
C++ mirror cpp/linalg.hpp) with g++, links it to liblinalg_b200.so and runs it on the GPU."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build(tmp_path):
    exe = str(tmp_path / "reference_suite")
    libdir = os.path.join(ROOT, "linear_algebra_b200")
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", os.path.join(ROOT, "tests", "reference_suite.cpp"), "-o", exe,
                    "-L" + libdir, "-llinalg_b200", "-Wl,-rpath," + libdir], check=True)
    return exe


def test_cpp_mirror_compiles(tmp_path):
    build(tmp_path)


@pytest.mark.gpu
def test_cpp_mirror_runs_the_reference_suite(tmp_path):
    r = subprocess.run([build(tmp_path)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "all reference tests passed" in r.stdout
