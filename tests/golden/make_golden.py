#!/usr/bin/env python
"""Generate tests/golden/model_*.json: model-level golden vectors produced by the UNMODIFIED
reference code (oracle/_ref, built from /root/reference by oracle/Makefile).

Run in the build container (needs oracle/_ref):  python tests/golden/make_golden.py
Each file holds a model state string, input points, and the reference's outputs
(coordinates, achievedPrecision, status bits) for groundToImage and imageToGround at two
precisions.  Doubles are stored as C99 hex strings so the fixtures are bit-exact.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref  # noqa: E402
from usgscsm_b200 import synth  # noqa: E402


def hexes(a):
    return [float(v).hex() for v in np.asarray(a, dtype=np.float64).reshape(-1)]


CASES = {
    "linescan_nac": lambda: synth.linescan_state(n_lines=2048, n_samples=1024, seed=101),
    "linescan_adjusted": lambda: synth.linescan_state(
        n_lines=2048, n_samples=1024, seed=102, distortion_type=0, dist_coeffs=(1e-4, -2e-7, 3e-10),
        params=(3.0, -2.0, 1.5, 0.02, -0.01, 0.005, 0.5, -0.3, 0.2, 0.05, 0.04, -0.03, 0.01, -0.02, 0.015, 0.05)),
    "pushframe_wac": lambda: synth.pushframe_state(n_framelets=64, framelet_height=16, n_samples=1024, seed=103),
    "frame_radial": lambda: synth.frame_state(n_lines=1024, n_samples=1024, seed=104),
    "frame_jitter": lambda: synth.frame_state(n_lines=256, n_samples=256, seed=105,
                                              jitter=((1e-2, -3e-2, 0.5), (-2e-2, 1e-2, -0.3), 1e-3)),
    "sar_orbital": lambda: synth.sar_state(n_lines=4000, n_samples=800, seed=106),
}


def main():
    for name, make in CASES.items():
        state = make()
        rm = ref.RefModel(state)
        nl, ns = rm.image_size()
        rng = np.random.default_rng(sum(map(ord, name)))
        n = 96
        line = rng.uniform(0, nl, n)
        samp = rng.uniform(0, ns, n)
        line[-8:] = rng.uniform(-2 * nl, 3 * nl, 8)   # a few pixels outside the image
        h = rng.uniform(-1000.0, 1000.0, n)
        h[-4:] = -1.0e7                                # rays that miss the body
        out = {"generator": "tests/golden/make_golden.py", "source": "oracle/_ref (reference @ ba32d13)",
               "state": state, "image_size": [nl, ns], "line": hexes(line), "samp": hexes(samp),
               "height": hexes(h), "cases": []}
        if name.startswith("linescan"):
            # the host-fitted projective seed is an INPUT of the hot path
            out["ls_derived"] = hexes(rm.ls_derived())
        for prec in (0.001, 1e-8):
            x, y, z, a, st = rm.image_to_ground(line, samp, h, prec)
            gx, gy, gz = x.copy(), y.copy(), z.copy()
            bad = ~(np.isfinite(gx) & np.isfinite(gy) & np.isfinite(gz))
            gx[bad], gy[bad], gz[bad] = 1.0e6, 2.0e6, -3.0e5  # some in-space point instead of NaN
            l, s, a2, st2 = rm.ground_to_image(gx, gy, gz, prec)
            out["cases"].append({
                "precision": prec,
                "i2g": {"x": hexes(x), "y": hexes(y), "z": hexes(z), "achieved": hexes(a),
                        "status": [int(v) for v in st]},
                "g2i_in": {"x": hexes(gx), "y": hexes(gy), "z": hexes(gz)},
                "g2i": {"line": hexes(l), "samp": hexes(s), "achieved": hexes(a2),
                        "status": [int(v) for v in st2]},
            })
        path = os.path.join(HERE, "model_%s.json" % name)
        with open(path, "w") as f:
            json.dump(out, f)
        print(name, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
