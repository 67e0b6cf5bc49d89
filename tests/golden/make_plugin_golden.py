#!/usr/bin/env python3
"""Generates tests/golden/plugin_goldens.json (run HERE, where /root/reference exists and oracle/_ref is built):

  for every case of tests/plugin_cases.py
    "golden"     the text of the reference's own golden output file (src/.../output/*.out) when the case has one
    "reference"  stdout of the UNMODIFIED reference program (oracle/_ref/bin/<prog>, reference CPU Mat/Vec types)
The GPU tests run the same binaries with `-dll_prepend libpetscb200plugin.so -mat_type aijb200 -vec_type b200` and
compare with both.
"""
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from plugin_cases import CASES, BIN  # noqa: E402

REF = "/root/reference"


def main():
    out = {}
    for name, c in CASES.items():
        exe = os.path.join(BIN, c["prog"])
        r = subprocess.run([exe] + c["args"] + c.get("ref_args", []), capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, (name, r.stdout[-500:], r.stderr[-500:])
        e = {"reference": r.stdout}
        if c.get("golden"):
            e["golden"] = open(os.path.join(REF, c["golden"])).read()
            e["golden_source"] = c["golden"]
        out[name] = e
        print("%-28s %5d bytes of reference output%s" % (name, len(r.stdout), ", golden " + c["golden"] if c.get("golden") else ""))
    json.dump(out, open(os.path.join(HERE, "plugin_goldens.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
