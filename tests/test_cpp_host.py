"""The C++ host layer (include/beifong_b200/host.hpp): built with g++ against libbeifong_b200.so and run as a program.

CPU: the error behaviour that mirrors the reference's Throw / NotImplementedError sites, and the loud failure of commit()
without a CUDA device.  GPU: Integrator::receive through the C++ classes against the oracle, and cancel()."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "beifong_b200")
BUILD = os.path.join(ROOT, "tests", "cpp", "_build")


def build_program():
    os.makedirs(BUILD, exist_ok=True)
    exe = os.path.join(BUILD, "host_radar")
    src = os.path.join(ROOT, "tests", "cpp", "host_radar.cpp")
    hdr = os.path.join(ROOT, "include", "beifong_b200", "host.hpp")
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), src, "-o", exe,
                               "-L" + LIBDIR, "-lbeifong_b200", "-ldl", "-lpthread", "-Wl,-rpath," + LIBDIR])
    return exe


def test_cpp_host_layer_errors(native):
    out = subprocess.run([build_program(), "--host-only"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "0 failure(s)" in out.stdout


@pytest.mark.gpu
def test_cpp_host_layer_receive_vs_oracle(native):
    oracle = os.path.join(ROOT, "oracle", "liboracle.so")
    out = subprocess.run([build_program(), oracle], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "0 failure(s)" in out.stdout and "cancel: stopped after" in out.stdout
