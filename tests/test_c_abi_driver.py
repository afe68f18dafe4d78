"""The boundary from plain C: tests/c/abi_driver.c includes include/ftf.h, is compiled as C99 with -pedantic -Werror and
links libftf_b200.so -- no C++ type, no torch, no Python between the caller and the library, which is the position the
JNI shim (jni/ftf_jni.cc) is in.  Without a GPU it exercises the host codecs and checks that create fails loudly; on a
B200 it drives create -> run -> submit/poll -> flush -> destroy against the known answers of RegressITCase.scala:35."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(ftf):
    out_dir = os.path.join(ROOT, "tests", "c", "build")
    os.makedirs(out_dir, exist_ok=True)
    exe = os.path.join(out_dir, "abi_driver")
    lib_dir = os.path.join(ROOT, "flink-tensorflow_b200")
    cmd = ["gcc", "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "c", "abi_driver.c"), "-o", exe, "-L", lib_dir, "-l:libftf_b200.so", "-Wl,-rpath," + lib_dir]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert p.returncode == 0, p.stdout
    return exe


def test_c99_driver_compiles_and_runs_the_host_side_without_a_device(ftf, model_dir):
    import torch
    if torch.cuda.is_available():
        pytest.skip("the device variant of this test runs here")
    p = subprocess.run([_build(ftf), model_dir, "--no-device"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert p.returncode == 0 and "abi_driver ok" in p.stdout, p.stdout


@pytest.mark.gpu
def test_c99_driver_drives_the_whole_lifecycle_on_the_gpu(ftf, model_dir):
    p = subprocess.run([_build(ftf), model_dir], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert p.returncode == 0 and "abi_driver ok" in p.stdout, p.stdout
