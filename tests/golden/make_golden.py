"""Generates tests/golden/golden.json: the digests the CPU oracle produces for every scenario in tests/scenarios.py.

The reference itself cannot be built or imported in this image (C++23 + Boost/Taskflow), so these
fixtures are outputs of the oracle restatement — pinned by oracle/kat_tests.cpp against the reference's own doctest
expectations — not of the reference binary.  Run:  python tests/golden/make_golden.py"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

import scenarios  # noqa: E402
from heliosengine_b200 import capi  # noqa: E402

if __name__ == "__main__":
    o = capi.load_oracle()
    out = {name: fn(o) for name, fn in scenarios.SCENARIOS.items()}
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print("wrote", len(out), "scenarios")
