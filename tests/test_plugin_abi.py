"""The dynamic-plugin boundary (SURVEY.md §8b seams B1 / B2; reference src/app/include/helios/app/plugin.hpp:14-62,
dynamic_plugin.hpp:27,34-36,84-88,399-445,541-586).

* tests/cpp/plugin_abi_host.cpp is a host that holds ONLY the reference's declaration of `Plugin`: it dlopens
  libhelios_gpu_plugin.so, resolves the two exported symbols, drives the five virtuals through the base pointer and deletes
  the plugin before dlclose — what DynamicPlugin::Load / Unload do.
* csrc/gpu_plugin_reference_headers.cpp — the plugin written against the reference's real headers — is syntax-checked
  against /root/reference as far as g++-13 gets (the reference needs GCC 14+ and Boost / Taskflow headers this image lacks):
  no diagnostic may point into our file.  Skipped where /root/reference does not exist (the GPU box)."""
import glob
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLUGIN = os.path.join(ROOT, "heliosengine_b200", "libhelios_gpu_plugin.so")
EXE = os.path.join(ROOT, "tests", "cpp", "plugin_abi_test")
VARIANT = os.path.join(ROOT, "heliosengine_b200", "csrc", "gpu_plugin_reference_headers.cpp")


@pytest.fixture(scope="module")
def abi_exe():
    subprocess.check_call(["make", "-s", "-j8", "-C", os.path.join(ROOT, "heliosengine_b200", "csrc")])
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "tests", "cpp"), "plugin_abi_test"])
    return EXE


def test_plugin_exports_exactly_the_reference_symbols(abi_exe):
    out = subprocess.run(["nm", "-D", "--defined-only", PLUGIN], capture_output=True, text=True).stdout
    syms = {l.split()[-1] for l in out.splitlines() if l.strip() and " T " in l}
    assert syms == {"helios_create_plugin", "helios_plugin_type_id"}, syms


def test_plugin_virtuals_through_the_reference_declaration_without_gpu(abi_exe):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present: the gpu variant covers this")
    r = subprocess.run([abi_exe, "nogpu", PLUGIN], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "0 failed" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_plugin_all_five_virtuals_on_gpu(abi_exe):
    r = subprocess.run([abi_exe, "gpu", PLUGIN], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "0 failed" in r.stdout, r.stdout + r.stderr


def test_real_header_variant_has_no_diagnostic_of_its_own():
    inc = sorted(glob.glob("/root/reference/src/*/include"))
    if not inc:
        pytest.skip("/root/reference is not present on this machine")
    cmd = ["g++", "-std=c++23", "-fsyntax-only", "-fmax-errors=50"] + [f"-I{d}" for d in inc] + [VARIANT]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    errors = re.findall(r"^(\S+?):(\d+):\d+: (?:fatal )?error: (.*)$", r.stderr, flags=re.M)
    ours = [e for e in errors if os.path.basename(e[0]) == os.path.basename(VARIANT)]
    assert not ours, ours
    # the reference itself does not get through g++-13 in this image: the first stop is a missing third-party header
    # (INTEGRATION.md §1 records it); if that ever changes, this test should start compiling the variant for real
    assert errors and errors[0][0].startswith("/root/reference/"), r.stderr[:2000]
