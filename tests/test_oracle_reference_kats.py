"""The reference's own known-answer tests for the hot path, re-expressed against the CPU oracle (oracle/kat_tests.cpp).

This is what pins the oracle: every check cites the doctest case under /root/reference/src/ecs/tests it restates.  The
reference itself cannot be compiled in this image, so there is no oracle/_ref binary to compare with."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_oracle_passes_reference_kats():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "kat"])
    r = subprocess.run([os.path.join(ROOT, "oracle", "kat_tests")], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 failed" in r.stdout
    n = int(r.stdout.split("oracle KATs:")[1].split("checks")[0])
    assert n >= 150
