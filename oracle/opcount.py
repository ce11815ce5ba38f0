#!/usr/bin/env python
"""Exact FP64 operation counts of the reference's per-point math, per BASELINE.json config
(TEST / MEASUREMENT INFRASTRUCTURE; SURVEY.md section 8(d): "replace these estimates by exact
counts from an op-counting build of the oracle math").

oracle/opcount.c single-steps the unmodified reference binary over a sample of each workload
and histograms the executed instruction addresses; this script maps them to mnemonics with
objdump and applies the counting convention of SURVEY.md 8(d):

    add / sub / mul / div / sqrt / compare / min / max / abs (andpd, andnpd) = 1 flop each
    (packed SSE2 forms = 2), a libm sin / cos call = 1, sincos = 2, other libm calls = 1,
    negation by xorpd = 0, moves / conversions / integer work = 0.

The reference is built without -march (no FMA), so there are no fused operations to split.
Per config it reports flops per point, the same split by function, and -- for the root
searches -- a linear model  flops/pt = per_evaluation x evaluations/pt + fixed  fitted from two
precisions, which bench.py applies to the evaluation count the kernel measures on the full
workload.  Output: profiles/r2_opcount.json.

    python oracle/opcount.py [--points 192]
"""
import argparse
import collections
import json
import os
import re
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
LIB = os.path.join(ROOT, "oracle", "_ref", "libusgscsm_ref.so")
TOOL = os.path.join(ROOT, "oracle", "_ref", "opcount")

ONE = {"addsd", "subsd", "mulsd", "divsd", "sqrtsd", "comisd", "ucomisd", "maxsd", "minsd",
       "cmpsd", "cmpltsd", "cmplesd", "cmpeqsd", "cmpneqsd", "cmpnltsd", "cmpnlesd", "cmpunordsd",
       "cmpordsd", "andpd", "andnpd"}
TWO = {"addpd", "subpd", "mulpd", "divpd", "sqrtpd", "maxpd", "minpd", "haddpd", "cmppd",
       "cmpltpd", "cmplepd", "cmpeqpd", "cmpneqpd", "cmpnltpd", "cmpnlepd", "addsubpd"}
KIND = {}
for m in ("addsd", "addpd", "subsd", "subpd", "haddpd", "addsubpd"):
    KIND[m] = "add_sub"
for m in ("mulsd", "mulpd"):
    KIND[m] = "mul"
for m in ("divsd", "divpd"):
    KIND[m] = "div"
for m in ("sqrtsd", "sqrtpd"):
    KIND[m] = "sqrt"
for m in ("andpd", "andnpd"):
    KIND[m] = "abs"


def disassembly():
    """address -> (mnemonic, enclosing symbol) for the whole library."""
    out = subprocess.run(["objdump", "-d", "-C", "--no-show-raw-insn", LIB], check=True,
                         capture_output=True, text=True).stdout
    table, sym = {}, "?"
    global ENTRY
    ENTRY = {}
    head = re.compile(r"^([0-9a-f]+) <(.*)>:$")
    ins = re.compile(r"^\s*([0-9a-f]+):\s+(\S+)")
    for ln in out.splitlines():
        m = head.match(ln)
        if m:
            sym = m.group(2)
            ENTRY[int(m.group(1), 16)] = sym
            continue
        m = ins.match(ln)
        if m:
            table[int(m.group(1), 16)] = (m.group(2), sym)
    return table


def short(sym):
    s = re.sub(r"\(.*", "", sym)
    s = s.replace("UsgsAstro", "").replace("SensorModel", "")
    return s.split(" ")[-1]


def count(state, direction, pts, prec, dis):
    """Run the op counter over `pts` (3 arrays); returns the per-point breakdown."""
    n = len(pts[0])
    with tempfile.TemporaryDirectory() as d:
        sp, pp, op = (os.path.join(d, f) for f in ("state.json", "pts.bin", "out.txt"))
        with open(sp, "w") as f:
            f.write(state)
        np.concatenate([np.asarray(p, dtype=np.float64) for p in pts]).tofile(pp)
        subprocess.run([TOOL, LIB, sp, direction, pp, str(n), repr(prec), op], check=True)
        lines = open(op).read().splitlines()
    flops = collections.Counter()
    kinds = collections.Counter()
    by_fn = collections.Counter()
    calls = collections.Counter()
    instr = 0
    entry = collections.Counter()
    for ln in lines:
        if ln.startswith("#"):
            header = ln
            continue
        a, b = ln.split()[:2] if not ln.startswith("libm") else (None, None)
        if ln.startswith("libm"):
            _, name, c = ln.split()
            c = int(c)
            w = 2 if name == "sincos" else 1
            calls[name] += c
            flops["libm"] += w * c
            kinds["libm_" + name] += c
            by_fn["libm::" + name] += w * c
            continue
        addr, c = int(a, 16), int(b)
        mn, sym = dis.get(addr, ("?", "?"))
        instr += c
        if addr in ENTRY and not ENTRY[addr].endswith("@plt"):
            entry[short(ENTRY[addr])] += c
        w = 1 if mn in ONE else (2 if mn in TWO else 0)
        if w:
            flops["fp64"] += w * c
            kinds[KIND.get(mn, "compare_min_max")] += w * c
            by_fn[short(sym)] += w * c
    # entries of the functions whose call counts define "evaluations" of a root search
    return {"points": n, "header": header,
            "flops_per_point": (flops["fp64"] + flops["libm"]) / n,
            "by_kind_per_point": {k: v / n for k, v in sorted(kinds.items())},
            "by_function_per_point": {k: v / n for k, v in by_fn.most_common(14)},
            "calls_per_point": {k: v / n for k, v in entry.most_common(12)},
            "library_instructions_per_point": instr / n}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=192)
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r2_opcount.json"))
    a = ap.parse_args()
    from oracle import ref
    from usgscsm_b200 import synth
    dis = disassembly()
    rng = np.random.default_rng(20261105)
    doc = {"how": "oracle/opcount.c: ptrace single-step of the unmodified reference binary "
                  "(oracle/_ref/libusgscsm_ref.so, -O3, no -march => no FMA), address histogram "
                  "x objdump mnemonics; convention of SURVEY.md 8(d)",
           "points_per_config": a.points, "configs": {}}

    def sample_grid(nl, ns, n):
        line = rng.integers(0, nl, n) + 0.5
        samp = rng.integers(0, ns, n) + 0.5
        return line, samp

    def run_model(name, state, nl, ns, i2g_too=True, precs=(0.001, 1e-8)):
        rm = ref.RefModel(state)
        line, samp = sample_grid(nl, ns, a.points)
        h = synth.synthetic_dem(line, samp)
        x, y, z, _, _ = rm.image_to_ground(line, samp, h)
        entry = {}
        if i2g_too:
            entry["image_to_ground"] = count(state, "i2g", (line, samp, h), 0.001, dis)
        for p in precs:
            entry["ground_to_image_prec_%g" % p] = count(state, "g2i", (x, y, z), p, dis)
        doc["configs"][name] = entry
        print(name, {k: round(v["flops_per_point"], 1) for k, v in entry.items()}, flush=True)
        return entry

    # C2: the bench workload (52224 x 5064 NAC-style line scan, LRO NAC distortion)
    run_model("C2_linescan_nac", synth.linescan_state(n_lines=52224, n_samples=5064, seed=20261104), 52224, 5064)
    # C3: push-frame and frame 4096 x 4096
    run_model("C3_pushframe_radial", synth.pushframe_state(n_framelets=256, framelet_height=16, n_samples=4096), 3584, 4096, precs=(0.001,))
    run_model("C3_frame_radial", synth.frame_state(n_lines=4096, n_samples=4096), 4096, 4096, precs=(0.001,))
    # C4: SAR
    run_model("C4_sar_orbital", synth.sar_state(), 64576, 2367, precs=(0.001,))
    # reference-held HiRISE state
    sup = open(os.path.join(ROOT, "tests", "golden", "ref_data", "gxp_model_file.sup"), "rb").read().decode("latin-1")
    run_model("gxp_hirise_sup", sup, 40000, 20000)
    # linear model of the line-scan search: flops/pt = per_evaluation x E + fixed, E = residual
    # evaluations of the secant search = computeDetectorView calls - 1 (the reference repeats the
    # view once after the search, LS:838-852)
    for name in ("C2_linescan_nac", "gxp_hirise_sup"):
        c = doc["configs"][name]
        lo, hi = c["ground_to_image_prec_0.001"], c["ground_to_image_prec_1e-08"]
        e = [v["calls_per_point"]["Ls::computeDetectorView"] - 1.0 for v in (lo, hi)]
        per = (hi["flops_per_point"] - lo["flops_per_point"]) / (e[1] - e[0])
        c["ground_to_image_model"] = {"flops_per_evaluation": per,
                                      "flops_fixed": lo["flops_per_point"] - per * e[0],
                                      "evaluations_per_point": {"prec_0.001": e[0], "prec_1e-08": e[1]}}
    with open(a.out, "w") as f:
        json.dump(doc, f, indent=1)


if __name__ == "__main__":
    main()
