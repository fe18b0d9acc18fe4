"""The C++ host layer (include/beifong_b200/host.hpp): built with g++ against libbeifong_b200.so and run as a program.

CPU: the error behaviour that mirrors the reference's Throw / NotImplementedError sites, and the loud failure of commit()
without a CUDA device.  GPU: Integrator::receive through the C++ classes against the oracle, and cancel()."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "beifong_b200")
BUILD = os.path.join(ROOT, "tests", "cpp", "_build")


def build_program(name="host_radar", extra=(), libs=()):
    os.makedirs(BUILD, exist_ok=True)
    exe = os.path.join(BUILD, name)
    src = os.path.join(ROOT, "tests", "cpp", name + ".cpp")
    hdr = os.path.join(ROOT, "include", "beifong_b200", "host.hpp")
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), *extra, src, "-o", exe,
                               "-L" + LIBDIR, "-lbeifong_b200", *libs, "-ldl", "-lpthread", "-Wl,-rpath," + LIBDIR])
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


@pytest.mark.gpu
def test_cpp_two_gpus_reduce_without_torch(native):
    """tests/cpp/multi_gpu.cpp: two devices from one C++ process through bf_multi_init / bf_multi_reduce (NCCL bound inside the
    library), the reduced profile against one device running the whole sample range.  Skips below two devices."""
    cuda = "/usr/local/cuda"
    exe = build_program("multi_gpu", extra=("-I" + os.path.join(cuda, "include"), "-L" + os.path.join(cuda, "lib64"), "-Wl,-rpath," + os.path.join(cuda, "lib64")), libs=("-lcudart",))
    out = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    if "SKIP" in out.stdout:
        pytest.skip(out.stdout.strip())
    assert "0 failure(s)" in out.stdout, out.stdout
