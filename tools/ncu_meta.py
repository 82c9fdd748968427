#!/usr/bin/env python
"""Writes <capture>.meta.json beside a tools/ncu_summary.py CSV: how many camera samples the captured launch rendered and
the fingerprint of the kernel sources it was built from (bench.py only trusts a capture whose fingerprint matches).
   python tools/ncu_meta.py profiles/r02_ncu_tile_cbox.csv <samples> [note]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import kernel_source_sha  # noqa: E402

path, samples = sys.argv[1], float(sys.argv[2])
json.dump({"samples": samples, "src_sha": kernel_source_sha(), "note": " ".join(sys.argv[3:])},
          open(path[:-4] + ".meta.json", "w"), indent=1)
print("wrote", path[:-4] + ".meta.json")
