"""bench.py's reference arm on the CPU: prints the contract's JSON line, runs none of the
product (the product's shared library is not even mapped into the process), and its config
equals the engine arm's for the same flags."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("gpus", [1, 2])
def test_reference_arm_line_and_no_product_library(ref, gpus):
    code = (
        "import runpy, sys\n"
        "sys.argv = ['bench.py', '--impl', 'reference', '--gpus', '%d', '--steps', '1', '--warmup', '0',"
        " '--lines', '1024', '--images', '4']\n"
        "runpy.run_path(%r, run_name='__main__')\n"
        "maps = open('/proc/self/maps').read()\n"
        "print('MAPPED', 'libusgscsm_b200.so' in maps, 'libusgscsm_ref.so' in maps)\n" % (gpus, os.path.join(ROOT, "bench.py")))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert lines[-1] == "MAPPED False True"
    doc = json.loads(lines[-2])
    assert doc["impl"] == "reference" and doc["n_gpus"] == gpus and doc["gpu_launches"] == 0
    assert doc["value"] > 0 and doc["unit"] == "Mpoints/s" and doc["cpu_baseline"]["kind"] == "reference"
    assert doc["e2e"]["h2d_bytes_per_step"] == 0 and doc["e2e"]["value"] == doc["value"]
    # the engine arm builds its config with the same function from the same flags
    sys.path.insert(0, ROOT)
    import bench
    a = type("A", (), {"gpus": gpus, "lines": 1024, "images": 4})()
    assert doc["config"] == bench.workload_config(a)
