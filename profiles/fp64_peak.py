#!/usr/bin/env python
"""Measured FP64 peaks of the box's GPU (the roofline denominators of bench.py; BASELINE.md
section 3): dependent-free DFMA / DADD / DMUL / DADD+DMUL streams on all SMs for >= 2 s each
(usgscsm_b200_measure_fp64, usgscsm_b200/csrc/fp64_peak.cu), with the SM clock sampled by
nvidia-smi during each run.  Writes one JSON document (default gpurun_out/r2_fp64_peak.json;
the committed copy is profiles/r2_fp64_peak.json).

    python profiles/fp64_peak.py [out.json] [seconds]
"""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "r2_fp64_peak.json")
    seconds = float(sys.argv[2]) if len(sys.argv) > 2 else 2.0
    import torch
    from bench import ClockSampler
    from usgscsm_b200 import _capi
    props = torch.cuda.get_device_properties(0)
    sms = props.multi_processor_count
    doc = {"gpu": props.name, "sms": sms, "seconds_per_op": seconds, "ops": {},
           "how": "usgscsm_b200_measure_fp64: 4 blocks x 256 threads per SM, 8 independent "
                  "accumulators per thread, CUDA events over repeated launches"}
    for op, name, flops in ((0, "dfma", 2), (1, "dadd", 1), (2, "dmul", 1), (3, "dadd_dmul_mix", 1)):
        ips, mhz = C.c_double(), C.c_double()
        s = ClockSampler(0)
        s.start()
        _capi.check(_capi.lib.usgscsm_b200_measure_fp64(0, op, seconds * 1e3, C.byref(ips), C.byref(mhz)))
        clocks = s.stop()
        sm_mhz = clocks.get("sm_mhz") or mhz.value
        doc["ops"][name] = {
            "thread_instructions_per_s": ips.value,
            "tflops": ips.value * flops / 1e12,
            "instructions_per_sm_per_clock": ips.value / (sms * sm_mhz * 1e6),
            "clocks": clocks,
        }
    doc["fp64_fma_peak_tflops"] = doc["ops"]["dfma"]["tflops"]
    doc["fp64_issue_peak_tinstr_per_s"] = doc["ops"]["dadd_dmul_mix"]["thread_instructions_per_s"] / 1e12
    os.makedirs(os.path.dirname(out), exist_ok=True)
    with open(out, "w") as f:
        json.dump(doc, f, indent=1)
    print(json.dumps(doc))


if __name__ == "__main__":
    main()
