"""The reference-shaped C++ host API (include/helios_b200/{ecs,app}.hpp) and the dynamic plugin library
(libhelios_gpu_plugin.so: helios_create_plugin / helios_plugin_type_id) — tests/cpp/shim_test.cpp."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLUGIN = os.path.join(ROOT, "heliosengine_b200", "libhelios_gpu_plugin.so")
EXE = os.path.join(ROOT, "tests", "cpp", "shim_test")


@pytest.fixture(scope="module")
def shim_exe():
    subprocess.check_call(["make", "-s", "-j8", "-C", os.path.join(ROOT, "heliosengine_b200", "csrc")])
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "tests", "cpp")])
    return EXE


def test_plugin_exports_reference_symbols(shim_exe):
    out = subprocess.run(["nm", "-D", "--defined-only", PLUGIN], capture_output=True, text=True).stdout
    syms = {l.split()[-1] for l in out.splitlines() if l.strip()}
    assert {"helios_create_plugin", "helios_plugin_type_id"} <= syms


def test_shim_compiles_and_fails_loudly_without_gpu(shim_exe):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    r = subprocess.run([shim_exe, "nogpu", PLUGIN], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 failed" in r.stdout


@pytest.mark.gpu
def test_shim_on_gpu(shim_exe):
    r = subprocess.run([shim_exe, "gpu", PLUGIN], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 failed" in r.stdout
