"""The C-ABI library loads without a GPU and exports every symbol include/helios_gpu.h declares (no compute calls)."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "helios_gpu.h")
LIB = os.path.join(ROOT, "heliosengine_b200", "libhelios_gpu.so")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"HELIOS_GPU_API\s+[\w\s\*]+?\b(helios_gpu_\w+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(LIB):
        subprocess.check_call(["make", "-j8", "-C", os.path.join(ROOT, "heliosengine_b200", "csrc")])
    return ctypes.CDLL(LIB)


def test_header_declares_the_boundary():
    syms = declared_symbols()
    assert len(syms) >= 45
    for must in ("helios_gpu_world_create", "helios_gpu_query_create", "helios_gpu_commands_enqueue",
                 "helios_gpu_schedule_run_and_wait", "helios_gpu_message_register", "helios_gpu_world_update"):
        assert must in syms


def test_library_exports_every_declared_symbol(lib):
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, missing


def test_no_oracle_or_cpu_fallback_in_product():
    """The product sources never reference oracle/ (a CPU route would void parity claims)."""
    pkg = os.path.join(ROOT, "heliosengine_b200")
    for dirpath, _, files in os.walk(os.path.join(pkg, "csrc")):
        for f in files:
            if f.endswith((".cu", ".cuh", ".hpp", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "helios_oracle" not in text and "oracle/" not in text, f
    # the shared library has no dependency on the oracle
    out = subprocess.run(["ldd", LIB], capture_output=True, text=True).stdout
    assert "oracle" not in out


def test_world_create_fails_loudly_without_a_device(lib):
    """No GPU here: world_create must return ERR_CUDA (-2), never a silent CPU world."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    h = ctypes.c_void_p()
    lib.helios_gpu_world_create.restype = ctypes.c_int32
    st = lib.helios_gpu_world_create(None, ctypes.byref(h))
    assert st == -2 and not h.value
