#!/usr/bin/env python3
"""Collect the reference's own golden outputs for its np > 1 test targets (the files tests/test_mpi_oracle.py::GOLDEN_RUNS
names, under /root/reference/src/*/examples/*/output/) into tests/golden/ref_mpi_goldens.json, so that the tests can
run where /root/reference is absent (the GPU box)."""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from test_mpi_oracle import GOLDEN_RUNS  # noqa: E402

out = {}
for _, _, _, rel, _ in GOLDEN_RUNS:
    out[rel] = open(os.path.join("/root/reference", rel)).read()
json.dump(out, open(os.path.join(HERE, "ref_mpi_goldens.json"), "w"), indent=1, sort_keys=True)
print("wrote %d goldens" % len(out))
