"""The host driver (csrc/mcc.cu: build_structure()'s index tables, the frame containers, the LM loop's bookkeeping)
under AddressSanitizer + UBSan.  libmcc_asan.so is the product's sources with instrumented host
code (csrc/Makefile, target asan); the test drivers are the facade programs of tests/cpp, linked against it.
compute-sanitizer is closed on the GPU pool, so this is the memory checker of the host side."""
import os
import subprocess

import pytest

from conftest import ROOT
from test_cpp_facade import _write_rig_file

PKG = os.path.join(ROOT, "multicam_calibration_b200")
GXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
BAD = ("AddressSanitizer", "runtime error:", "LeakSanitizer")


def _build(tmp_path, src):
    exe = tmp_path / (os.path.splitext(src)[0] + "_asan")
    subprocess.check_call([GXX, "-O1", "-g", "-std=c++17", "-fsanitize=address,undefined", "-fno-omit-frame-pointer", "-I", os.path.join(ROOT, "include"),
                           "-o", str(exe), os.path.join(ROOT, "tests", "cpp", src), "-L", PKG, "-lmcc_asan", f"-Wl,-rpath,{PKG}"])
    return exe


def test_host_logic_under_sanitizers(built, tmp_path):
    """tests/cpp/test_facade.cpp (containers, YAML / corners I/O, error paths; no GPU) with leak detection on"""
    exe = _build(tmp_path, "test_facade.cpp")
    r = subprocess.run([str(exe), str(tmp_path)], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, UBSAN_OPTIONS="print_stacktrace=1:halt_on_error=1"))
    out = r.stdout + r.stderr
    assert r.returncode == 0 and "FACADE_OK" in r.stdout and not any(b in out for b in BAD), out[-4000:]


@pytest.mark.gpu
@pytest.mark.parametrize("devices", ["-", "0,1"])
def test_solve_under_sanitizers(built, tmp_path, devices):
    """setCameras / addPoints / runCalibration / testCalibration on a GPU with the host code instrumented: one device,
    and one handle over two devices (frame scatter, member threads, in-process rendezvous)"""
    import torch
    from multicam_calibration_b200 import synth
    if devices != "-" and torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    rig = synth.make_synthetic_rig(3, 30, 20260111, detect_prob=0.85)
    _write_rig_file(rig, tmp_path / "rig.txt")
    exe = _build(tmp_path, "facade_gpu.cpp")
    # CUDA maps memory where ASan keeps its shadow gap; its driver-side allocations are not this library's leaks
    env = dict(os.environ, ASAN_OPTIONS="protect_shadow_gap=0:detect_leaks=0", UBSAN_OPTIONS="print_stacktrace=1:halt_on_error=1")
    r = subprocess.run([str(exe), str(tmp_path / "rig.txt"), devices, "run"], capture_output=True, text=True, timeout=600, env=env)
    out = r.stdout + r.stderr
    assert r.returncode == 0 and "FACADE_GPU_OK" in r.stdout and not any(b in out for b in BAD), out[-4000:]
