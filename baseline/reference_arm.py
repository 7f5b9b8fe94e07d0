"""Reference arm of bench.py (`--impl reference`).

Runs the UNMODIFIED reference code installed under ``baseline/_ref`` (git-ignored, built
once from /root/reference by ``baseline/build_ref.sh``).  See DESIGN.md "Reference arm".
"""
from __future__ import annotations

import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))


def run_reference(args) -> int:
    lib = os.path.join(HERE, "_ref", "libref_harness.so")
    if not os.path.exists(lib):
        print(json.dumps({"impl": "reference",
                          "unavailable": "baseline/_ref is not built (run baseline/build_ref.sh)"}))
        return 0
    from .ref_driver import run
    return run(args, lib)
