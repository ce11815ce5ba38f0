"""A/B runs of bench.py under different tuning-knob environments: each
setting in its own process (the knobs are read once), device-resident rates only.
    python profiles/ab_env.py VAR=v1,v2,... [VAR2=...]  [-- extra bench flags]"""
import itertools
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
args = sys.argv[1:]
extra = []
if "--" in args:
    k = args.index("--")
    args, extra = args[:k], args[k + 1:]
names = [a.split("=")[0] for a in args]
values = [a.split("=")[1].split(",") for a in args]
for combo in itertools.product(*values):
    env = dict(os.environ, **dict(zip(names, combo)))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--no-cpu", "--no-others", "--no-e2e",
                        "--steps", "3", "--warmup", "3"] + extra, env=env, capture_output=True, text=True)
    try:
        d = json.loads(r.stdout.strip().splitlines()[-1])
        print(dict(zip(names, combo)), "g2i %.0f Mpt/s (%.2f ms)  i2g grid %.0f (%.2f ms)  i2g explicit %.0f  g2i dem %.0f  g2i 1e-8 %.0f"
              % (d["value"], d["roofline"]["kernel_ms"], d["image_to_ground"]["value"], d["image_to_ground"]["kernel_ms"],
                 d["image_to_ground_explicit"]["value"], d["ground_to_image_dem_grid"]["value"],
                 d["ground_to_image_prec_1e-8"]["value"]), flush=True)
    except Exception as e:
        print(dict(zip(names, combo)), "FAILED", e, r.stderr[-500:], flush=True)
