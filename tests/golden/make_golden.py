#!/usr/bin/env python
"""Regenerates tests/golden/ from the REFERENCE BINARY (oracle/_ref/compress_ref).

Run in the build container only (needs /root/reference to have been compiled by
`make -C oracle ref`):   python tests/golden/make_golden.py

For every named input of SURVEY.md §8c it writes the input recipe, runs the
reference's `compress c fl-cpu` on it, and records header / sha256 / (for small
cases) the whole compressed file.  It also round-trips each through the
reference's `d fl-cpu`.  The RL entries are the worked examples of
IMPLEMENTATION-PLAN.md (the reference has no RL code to run).
"""
import hashlib
import json
import os
import shutil
import struct
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
from golden_inputs import INPUTS, make_input  # noqa: E402

REF_CLI = os.path.join(ROOT, "oracle", "_ref", "compress_ref")
REF_BMP = "/root/reference/example_input/sample_1280×853.bmp"
SMALL = 16384  # compressed files up to this size are stored whole


def run_ref(args):
    subprocess.run([REF_CLI] + args, check=True, stdout=subprocess.DEVNULL)


def main():
    assert os.path.exists(REF_CLI), "build oracle/_ref first: make -C oracle ref"
    bmp_dst = os.path.join(HERE, "sample_1280x853.bmp")
    if os.path.exists(REF_BMP) and not os.path.exists(bmp_dst):
        shutil.copyfile(REF_BMP, bmp_dst)  # data fixture (not source); config[0] of BASELINE.json
    entries = []
    with tempfile.TemporaryDirectory() as td:
        for name in INPUTS:
            data = make_input(name, HERE)
            src, comp, back = (os.path.join(td, x) for x in ("in", "out.fl", "back"))
            with open(src, "wb") as f:
                f.write(data)
            run_ref(["c", "fl-cpu", src, comp])
            run_ref(["d", "fl-cpu", comp, back])
            blob = open(comp, "rb").read()
            rt = open(back, "rb").read()
            assert rt == data, name
            n, nb, nv = struct.unpack_from("<QQQ", blob, 0)
            e = {
                "name": name,
                "input_sha256": hashlib.sha256(data).hexdigest(),
                "header": [n, nb, nv],
                "file_len": len(blob),
                "sha256": hashlib.sha256(blob).hexdigest(),
            }
            if len(blob) <= SMALL:
                fn = name + ".fl"
                with open(os.path.join(HERE, fn), "wb") as f:
                    f.write(blob)
                e["file"] = fn
            entries.append(e)
            print(name, e["header"], e["file_len"], e["sha256"][:16])
    rl = [
        {"name": "plan_encode", "cite": "IMPLEMENTATION-PLAN.md:87-89",
         "input": [5, 5, 8, 8, 8, 7, 7, 7, 7, 3, 4, 4, 4], "counts": [2, 3, 4, 1, 3], "values": [5, 8, 7, 3, 4]},
        {"name": "plan_decode", "cite": "IMPLEMENTATION-PLAN.md:158-160",
         "input": [8, 8, 8, 9, 9, 2, 4, 4], "counts": [3, 2, 1, 2], "values": [8, 9, 2, 4]},
        {"name": "plan_split_256", "cite": "IMPLEMENTATION-PLAN.md:125",
         "input": [7] * 256, "counts": [255, 1], "values": [7, 7]},
    ]
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump({"generator": "tests/golden/make_golden.py", "reference_cli": "compress c fl-cpu",
                   "fl": entries, "rl": rl}, f, indent=1)


if __name__ == "__main__":
    main()
