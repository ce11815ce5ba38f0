"""Loader of tests/golden/model_*.json (written by tests/golden/make_golden.py)."""
import glob
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def unhex(v):
    return np.array([float.fromhex(s) for s in v], dtype=np.float64)


def golden_files():
    return sorted(glob.glob(os.path.join(HERE, "golden", "model_*.json")))


def load(path):
    g = json.load(open(path))
    for k in ("line", "samp", "height"):
        g[k] = unhex(g[k])
    if "ls_derived" in g:
        g["ls_derived"] = unhex(g["ls_derived"])
    for c in g["cases"]:
        for blk in ("i2g", "g2i_in", "g2i"):
            for k, v in c[blk].items():
                c[blk][k] = np.array(v, dtype=np.uint32) if k == "status" else unhex(v)
    return g
