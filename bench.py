#!/usr/bin/env python
"""bench.py -- line-scan groundToImage / imageToGround throughput on B200 (BASELINE.json metric).

Workload (config C2 of BASELINE.json): a synthetic orbital NAC-style line-scan image
(lroLrocNac distortion), 5064 samples x 52224 lines per GPU, dense grid orthorectification:
imageToGround of every pixel centre at a synthetic-DEM height, then groundToImage of every
resulting ground point.  A "step" is one groundToImage pass over the whole grid (264.5 M
points per GPU); imageToGround is timed beside it and reported under "image_to_ground".
At N > 1 the image is N times taller and every rank takes one row band (weak scaling, no
data-path collective).

  python bench.py [--gpus N] [--steps K] [--warmup W]          # this engine
  python bench.py --impl reference [...]                        # the reference's CPU code
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mpoints/s groundToImage & imageToGround (line-scan) at 1/2/4/8 B200 vs host CPU"
N_SAMPLES = 5064
N_LINES_PER_GPU = 52224
PREC = 0.001
# algorithmic HBM bytes per point: x,y,z in; line,samp + status out
G2I_BYTES_PER_POINT = 24 + 16 + 4
# dense-grid imageToGround: height in; x,y,z + status out
I2G_BYTES_PER_POINT = 8 + 24 + 4
# algorithmic FP64 flops (SURVEY.md §8d counting convention): per evaluation of the detector
# view on the no-adjustment path with LRO-NAC distortion, and per point outside the loop
FLOPS_PER_EVALUATION = 450.0
FLOPS_FIXED = 40.0


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--lines", type=int, default=N_LINES_PER_GPU,
                    help="lines per GPU (default: the full C2 image)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-others", action="store_true",
                    help="skip the push-frame / frame / SAR side measurements (configs C3, C4)")
    return ap.parse_args()


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.rows.append([c.strip() for c in ln.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(self.NAMES, r[2:]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def workload_state(image, lines):
    """Image `image` of the job: a C2-shaped line-scan state (5064 x `lines`), its own orbit
    phase and attitude wobble (seed 20261104 + image)."""
    from usgscsm_b200 import synth
    return synth.linescan_state(n_lines=lines, n_samples=N_SAMPLES, seed=20261104 + image)


def peak_hbm():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def sample_pixels(n_lines_total, target_points):
    """Bounded sample of the dense grid: every k-th line, all samples (pixel centres)."""
    import numpy as np
    lines_wanted = max(1, min(n_lines_total, int(target_points // N_SAMPLES)))
    k = max(1, n_lines_total // lines_wanted)
    rows = np.arange(0, n_lines_total, k, dtype=np.float64)[:lines_wanted] + 0.5
    line = np.repeat(rows, N_SAMPLES)
    samp = np.tile(np.arange(N_SAMPLES, dtype=np.float64) + 0.5, rows.size)
    return line, samp, "every %d-th line of the grid, all %d samples: %d points" % (k, N_SAMPLES, line.size)


def cpu_reference_rate(state, n_lines_total, seconds, direction="g2i", proj=None, cores=None):
    """Reference CPU code (oracle/_ref, OpenMP over points, all host threads) on a bounded
    sample sized for about `seconds` of work.  Returns (Mpoints/s, cores, kind, sample, fn)
    where fn() re-times one pass over the sample."""
    import numpy as np
    from oracle import ref
    from usgscsm_b200 import synth
    rm = ref.RefModel(state)
    if proj is not None:
        rm.ls_set_projective(True, proj)
    # every host core this process may run on -- not OMP_NUM_THREADS, which torchrun sets to 1
    if cores is None:
        cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)

    def run(line, samp):
        h = synth.synthetic_dem(line, samp)
        x, y, z, _, _ = rm.image_to_ground(line, samp, h, PREC, threads=cores)
        if direction == "i2g":
            t = time.perf_counter()
            rm.image_to_ground(line, samp, h, PREC, threads=cores)
            return time.perf_counter() - t, None
        t = time.perf_counter()
        rm.ground_to_image(x, y, z, PREC, threads=cores)
        return time.perf_counter() - t, (x, y, z)

    line, samp, _ = sample_pixels(n_lines_total, 40 * N_SAMPLES)
    dt, _ = run(line, samp)
    rate = line.size / dt
    line, samp, desc = sample_pixels(n_lines_total, rate * seconds)
    dt, pts = run(line, samp)

    def again():
        t = time.perf_counter()
        if direction == "i2g":
            rm.image_to_ground(line, samp, synth.synthetic_dem(line, samp), PREC, threads=cores)
        else:
            rm.ground_to_image(pts[0], pts[1], pts[2], PREC, threads=cores)
        return time.perf_counter() - t, line.size

    return line.size / dt / 1e6, cores, "reference", desc, again


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import ref
    if not ref.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libusgscsm_ref.so not built"}))
        return
    # every image of the job has the C2 shape and costs the same per point: image 0 is sampled
    state = workload_state(0, args.lines)
    n_lines_total = args.lines
    # each step is one pass over a sample sized for ~4 s of host work
    _, cores, kind, desc, again = cpu_reference_rate(state, n_lines_total, 4.0)
    for _ in range(args.warmup):
        again()
    t_total, n_total = 0.0, 0
    for _ in range(args.steps):
        dt, n = again()
        t_total += dt
        n_total += n
    value = n_total / t_total / 1e6
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "Mpoints/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t_total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args),
        "cpu_baseline": {"value": value, "unit": "Mpoints/s", "cores": cores, "kind": kind,
                         "sample": desc + " per step, groundToImage"},
        "e2e": {"value": value, "unit": "Mpoints/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def workload_config(args):
    return {"workload": "C2 orbital line-scan (lroLrocNac distortion) %dx%d dense grid, "
                        "imageToGround at synthetic-DEM heights then groundToImage; value = "
                        "groundToImage; at N GPUs the job is N such images (C5 style)"
                        % (N_SAMPLES, args.lines),
            "points_per_gpu": N_SAMPLES * args.lines, "desired_precision": PREC,
            "sharding": ("%d images, each split into %d row bands, rank r takes band r of every "
                         "image; no data-path collective" % (args.gpus, args.gpus))
            if args.gpus > 1 else "single GPU",
            "l2": "inputs (6.3 GB per GPU) larger than L2; no flush needed"}


def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import usgscsm_b200 as ub
    from usgscsm_b200 import _capi, synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the engine has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa_node = None
    if world > 1:
        from usgscsm_b200 import sharding
        if os.environ.get("USGSCSM_B200_NUMA_BIND", "1") != "0":
            numa_node = sharding.bind_to_gpu_numa_node(local)
        # stdout carries ONE JSON line: keep NCCL's version banner (NCCL_DEBUG=VERSION) off it
        if os.environ.get("NCCL_DEBUG", "").upper() not in ("INFO", "TRACE"):
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # The job: `world` images of the C2 shape; every image is cut into `world` row bands and
    # this rank takes band `rank` of each (world == 1: the whole single image).  Per-GPU work
    # is the same at every N (weak scaling) and every image keeps the C2 table sizes.
    from usgscsm_b200 import sharding
    lines = args.lines
    first, band = sharding.band_for_rank(lines, rank, world)
    seg = band * N_SAMPLES                  # points of one band
    n = world * seg                         # points this rank processes per step
    n_job = world * lines * N_SAMPLES       # points of the whole job per step
    states = [workload_state(i, lines) for i in range(world)]
    plugin = ub.UsgsAstroPlugin(device=local)
    models = [plugin.constructModelFromState(s) for s in states]
    model, state = models[0], states[0]
    line0 = first + 0.5

    # synthetic DEM heights of the band (the same band of every image)
    h = torch.empty(n, dtype=torch.float64, device=dev)
    cols = torch.arange(N_SAMPLES, device=dev, dtype=torch.float64) + 0.5
    cterm = torch.cos(2 * np.pi * cols / 1024.0)
    for r0 in range(0, band, 4096):
        r1 = min(band, r0 + 4096)
        rows = torch.arange(r0, r1, device=dev, dtype=torch.float64) + line0
        h[r0 * N_SAMPLES:r1 * N_SAMPLES] = (500.0 * torch.sin(2 * np.pi * rows / 4096.0)[:, None] * cterm[None, :]).reshape(-1)
    for i in range(1, world):
        h[i * seg:(i + 1) * seg] = h[:seg]

    x, y, z = (torch.empty(n, dtype=torch.float64, device=dev) for _ in range(3))
    st_g = torch.empty(n, dtype=torch.int32, device=dev)
    out_line, out_samp = (torch.empty(n, dtype=torch.float64, device=dev) for _ in range(2))
    st_i = torch.empty(n, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    h_lib = model._h

    def off(t, i, itemsize=8):
        return t.data_ptr() + i * seg * itemsize

    def i2g():
        for i, m in enumerate(models):
            _capi.check(_capi.lib.usgscsm_b200_image_to_ground_grid_dev(
                m._h, line0, 0.5, 1.0, 1.0, band, N_SAMPLES, off(h, i), 1, PREC, off(x, i),
                off(y, i), off(z, i), None, off(st_g, i, 4), stream))

    def g2i(prec=PREC):
        for i, m in enumerate(models):
            _capi.check(_capi.lib.usgscsm_b200_ground_to_image_dev(
                m._h, seg, off(x, i), off(y, i), off(z, i), prec, off(out_line, i),
                off(out_samp, i), None, off(st_i, i, 4), stream))

    def timed(fn, steps):
        """K launches bracketed by barrier + synchronize; per-launch CUDA events on the
        launching stream.  Returns (total ms max over ranks, mean per-launch ms on this rank)."""
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
              for _ in range(steps)]
        barrier()
        t0 = torch.cuda.Event(enable_timing=True)
        t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        for a, b in ev:
            a.record()
            fn()
            b.record()
        t1.record()
        torch.cuda.synchronize()
        total = t0.elapsed_time(t1)
        per = sum(a.elapsed_time(b) for a, b in ev) / steps
        barrier()
        return max_over_ranks(total), per

    i2g()
    torch.cuda.synchronize()
    for _ in range(args.warmup):
        g2i()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = _capi.lib.usgscsm_b200_kernel_launches()
    total_ms, kern_ms = timed(g2i, args.steps)
    launches = _capi.lib.usgscsm_b200_kernel_launches() - launches0
    clocks = sampler.stop() if rank == 0 else None
    value = n_job * args.steps / (total_ms * 1e-3) / 1e6

    # correctness of what was just timed: the round trip must close and no flag may be set
    band_rows = (torch.arange(band, device=dev, dtype=torch.float64) + line0)[None, :, None]
    err = max(float((out_line.view(world, band, N_SAMPLES) - band_rows).abs().max()),
              float((out_samp.view(world, band, N_SAMPLES) - cols[None, None, :]).abs().max()))
    ev_mask = _capi.ST_EVALS_MASK
    flags = int(((st_i & ~ev_mask) != 0).sum().item()) + int(((st_g & ~ev_mask) != 0).sum().item())
    # evaluations of the detector view per point, counted by the kernel itself (status bits 16-23)
    mean_evals = float(((st_i >> _capi.ST_EVALS_SHIFT) & 0xff).double().mean().item())

    # imageToGround beside it
    for _ in range(args.warmup):
        i2g()
    i2g_total_ms, i2g_kern_ms = timed(i2g, args.steps)
    i2g_value = n_job * args.steps / (i2g_total_ms * 1e-3) / 1e6

    peak, peak_src = peak_hbm()
    traffic, traffic_src = ncu_traffic(n)
    achieved = G2I_BYTES_PER_POINT * n / (kern_ms * 1e-3) / 1e9
    # FP64 roofline: algorithmic flops per point = 450 per detector-view
    # evaluation x measured evaluations + 40; peak = 64 FP64 FMA lanes per SM x 2 flop x SMs x
    # the SM clock sampled under load (ncu: sm__sass_thread_inst_executed_op_dfma_pred_on
    # peak_sustained = 64 per SM per cycle).
    props = torch.cuda.get_device_properties(dev)
    sm_mhz = (clocks or {}).get("sm_mhz") or (clocks or {}).get("sm_max_mhz") or 1965.0
    fp64_peak_tflops = 64 * 2 * props.multi_processor_count * sm_mhz * 1e6 / 1e12
    flops_per_point = FLOPS_PER_EVALUATION * mean_evals + FLOPS_FIXED
    fp64_achieved = flops_per_point * n / (kern_ms * 1e-3) / 1e12
    result = {
        "metric": METRIC, "value": value, "unit": "Mpoints/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": workload_config(args), "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "traffic_unit": "GB per launch",
                     "traffic_source": traffic_src, "peak_source": peak_src,
                     "kernel": "lsGroundToImageKernel", "kernel_ms": kern_ms,
                     "bytes_per_point": G2I_BYTES_PER_POINT,
                     "note": "the kernel is FP64-pipe bound, not HBM bound (DESIGN.md §4): fp64 "
                             "below is the roofline that binds",
                     "fp64": {"bound": "fp64", "achieved": fp64_achieved, "peak": fp64_peak_tflops,
                              "unit": "TFLOP/s", "frac": fp64_achieved / fp64_peak_tflops,
                              "flops_per_point": flops_per_point,
                              "evaluations_per_point": mean_evals,
                              "peak_source": "64 DFMA lanes/SM/clk x 2 x %d SMs x %.0f MHz (FMA "
                                             "counted as 2; the reference's arithmetic has no "
                                             "FMAs, so 0.5 is the ceiling for identical "
                                             "rounding)" % (props.multi_processor_count, sm_mhz)}},
        "image_to_ground": {"value": i2g_value, "unit": "Mpoints/s", "ms_per_step": i2g_total_ms / args.steps,
                            "kernel_ms": i2g_kern_ms,
                            "hbm_gbs": I2G_BYTES_PER_POINT * n / (i2g_kern_ms * 1e-3) / 1e9},
        "check": {"round_trip_max_px": err, "points_with_flags": flags},
    }
    ncu = ncu_fp64()
    if ncu:
        result["roofline"]["fp64"]["ncu"] = ncu
    if world > 1:
        result["config"]["numa_bind"] = "rank 0 bound to NUMA node %s" % numa_node

    # second column: desiredPrecision 1e-8, what the reference's own CLI test
    # uses (tests/CMakeLists.txt:30) -- more secant iterations per point
    def g2i_tight():
        g2i(1e-8)
    g2i_tight()
    tight_total_ms, tight_kern_ms = timed(g2i_tight, args.steps)
    result["ground_to_image_prec_1e-8"] = {
        "value": n_job * args.steps / (tight_total_ms * 1e-3) / 1e6, "unit": "Mpoints/s",
        "kernel_ms": tight_kern_ms,
        "evaluations_per_point": float(((st_i >> _capi.ST_EVALS_SHIFT) & 0xff).double().mean().item())}

    # imageToGround from explicit (line, samp) arrays (not the dense-grid tables): 52 B/pt
    if world == 1:
        ex_line = (torch.arange(lines, device=dev, dtype=torch.float64) + line0).repeat_interleave(N_SAMPLES)
        ex_samp = cols.repeat(lines)

        def i2g_explicit():
            _capi.check(_capi.lib.usgscsm_b200_image_to_ground_dev(
                h_lib, n, ex_line.data_ptr(), ex_samp.data_ptr(), h.data_ptr(), 1, PREC,
                out_line.data_ptr(), out_samp.data_ptr(), ex_line.data_ptr(), None, st_i.data_ptr(), stream))
        i2g_explicit()
        ex_total_ms, ex_kern_ms = timed(i2g_explicit, max(1, args.steps // 2))
        result["image_to_ground_explicit"] = {
            "value": n * max(1, args.steps // 2) / (ex_total_ms * 1e-3) / 1e6,
            "unit": "Mpoints/s", "kernel_ms": ex_kern_ms}
        del ex_line, ex_samp

    # the callers' fan-out around the path, line-scan model, first 64 Mi points
    if world == 1 and not args.no_others:
        result["callers"] = callers_side(model, x, y, z, min(n, 1 << 26), dev, band)

    if not args.no_e2e:
        result["e2e"] = e2e(args, models, x, y, z, seg, n_job, barrier, max_over_ranks)
        if world == 1:
            result["e2e_image_to_ground_grid"] = e2e_grid(args, model, h, lines, line0)
    if rank == 0 and world == 1 and not args.no_cpu:
        from oracle import ref
        if ref.available():
            v, cores, kind, desc, _ = cpu_reference_rate(state, lines, 15.0, proj=model.lsDerived()[5:19])
            result["cpu_baseline"] = {"value": v, "unit": "Mpoints/s", "cores": cores, "kind": kind,
                                      "sample": desc + ", groundToImage, OpenMP over points"}
            v1, _, _, desc1, _ = cpu_reference_rate(state, lines, 4.0, proj=model.lsDerived()[5:19], cores=1)
            result["cpu_baseline"]["single_thread"] = {"value": v1, "sample": desc1}
    if rank == 0 and world == 1 and not args.no_others:
        result["other_models"] = other_models(dev)
    if rank == 0:
        print(json.dumps(result))
    if world > 1:
        dist.destroy_process_group()


def ncu_fp64():
    """FP64 pipe utilisation of lsGroundToImageKernel from the committed ncu --set full capture
    (profiles/ncu_g2i.json): share of cycles the pipe is busy and FP64 thread-instructions
    issued per cycle against the 64 lanes x 148 SMs the chip can issue."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "ncu_g2i.json")
    try:
        with open(path) as f:
            p = json.load(f)
        return {"pipe_active_pct": p["fp64_pipe_active_pct"],
                "issue_frac": p["fp64_thread_inst_per_cycle"] / p["fp64_issue_peak_per_cycle"],
                "kernel_ms_under_ncu": p["duration_ms"], "source": "profiles/ncu_g2i.json"}
    except (OSError, KeyError, ValueError):
        return None


def ncu_traffic(n):
    """DRAM bytes of one lsGroundToImageKernel launch from the committed ncu --set full capture
    (profiles/ncu_g2i.json: dram__bytes_read.sum + dram__bytes_write.sum and the points of the
    captured launch), scaled to this launch's points.  None if the summary is absent."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "ncu_g2i.json")
    try:
        with open(path) as f:
            p = json.load(f)
        per_point = (p["dram_bytes_read"] + p["dram_bytes_write"]) / p["points"]
        return per_point * n / 1e9, "profiles/ncu_g2i.json (%s)" % p.get("capture", "ncu --set full")
    except (OSError, KeyError, ValueError):
        return None, None


def callers_side(model, x, y, z, m, dev, band):
    """Side measurements of the batched caller-level entry points on the C2 line-scan model:
    computeGroundPartials (4 searches per point in one kernel), computeSensorPartials (2
    searches + difference), computeAllSensorPartials (17 searches) and groundToImage with
    covariance (22 searches) on a quarter of the points ("all_" keys), getImageTime +
    getSensorPosition + getSensorVelocity, getIlluminationDirection, imageToRemoteImagingLocus, imageToProximateImagingLocus (3
    imageToGround + 2 elevations).  Device resident, median of 3 after 1 warm-up."""
    import torch
    from usgscsm_b200 import _capi
    lib, h = _capi.lib, model._h
    stream = torch.cuda.current_stream(dev).cuda_stream
    f64 = lambda k: [torch.empty(m, dtype=torch.float64, device=dev) for _ in range(k)]
    st = torch.empty(m, dtype=torch.int32, device=dev)
    part = torch.empty(6 * m, dtype=torch.float64, device=dev)
    o = f64(7)
    rows = torch.arange(band, device=dev, dtype=torch.float64) + 0.5
    line = rows.repeat_interleave(N_SAMPLES)[:m].contiguous()
    samp = (torch.arange(N_SAMPLES, device=dev, dtype=torch.float64) + 0.5).repeat(band)[:m].contiguous()
    p = lambda t: t.data_ptr()
    calls = {
        "ground_partials": lambda: lib.usgscsm_b200_ground_partials_dev(
            h, m, p(x), p(y), p(z), p(part), p(st), stream),
        "sensor_partials_index0": lambda: lib.usgscsm_b200_sensor_partials_dev(
            h, 0, m, p(x), p(y), p(z), None, None, PREC, p(o[0]), p(o[1]), p(o[2]), p(st), stream),
        "remote_imaging_locus": lambda: lib.usgscsm_b200_image_to_remote_imaging_locus_dev(
            h, m, p(line), p(samp), PREC, *[p(t) for t in o], p(st), stream),
        "proximate_imaging_locus": lambda: lib.usgscsm_b200_image_to_proximate_imaging_locus_dev(
            h, m, p(line), p(samp), p(x), p(y), p(z), PREC, *[p(t) for t in o], p(st), stream),
    }
    npar = lib.usgscsm_b200_model_num_parameters(h)
    m4 = m // 4     # 17 searches and 256 B of partials per point
    allp = torch.empty(m4 * npar * 2, dtype=torch.float64, device=dev)
    calls["all_sensor_partials_%d_params" % npar] = lambda: lib.usgscsm_b200_all_sensor_partials_dev(
        h, npar, None, m4, p(x), p(y), p(z), None, None, PREC, p(allp), p(o[0]), p(st), stream)
    gcov = torch.eye(3, dtype=torch.float64, device=dev).reshape(1, 9).repeat(m4, 1).contiguous()
    icov = torch.empty(m4 * 4, dtype=torch.float64, device=dev)
    calls["all_ground_to_image_covar"] = lambda: lib.usgscsm_b200_ground_to_image_covar_dev(
        h, m4, p(x), p(y), p(z), p(gcov), PREC, p(o[0]), p(o[1]), p(icov), p(o[2]), p(st), stream)
    calls["image_sensor_state"] = lambda: lib.usgscsm_b200_image_sensor_state_dev(
        h, m, p(line), p(samp), *[p(t) for t in o], p(st), stream)
    calls["illumination_direction"] = lambda: lib.usgscsm_b200_illumination_direction_dev(
        h, m, p(x), p(y), p(z), p(o[0]), p(o[1]), p(o[2]), p(st), stream)
    out = {"points": m, "unit": "Mpoints/s", "points_all_": m4}
    for name, fn in calls.items():
        _capi.check(fn())
        torch.cuda.synchronize()
        ms = []
        for _ in range(3):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            _capi.check(fn())
            b.record()
            torch.cuda.synchronize()
            ms.append(a.elapsed_time(b))
        out[name] = (m4 if name.startswith("all_") else m) / sorted(ms)[1] / 1e3
    return out


def other_models(dev):
    """Side measurements (not the headline): dense-grid imageToGround at synthetic-DEM heights
    then groundToImage of the result for the push-frame and frame models (config C3,
    4096 x 4096) and the SAR model (config C4 shape, 64576 x 2367 = 152.9 M points), device
    resident, outputs preallocated, CUDA events per launch, median of 5 after 2 warm-ups."""
    import numpy as np
    import torch
    import usgscsm_b200 as ub
    from usgscsm_b200 import _capi, synth
    out = {}
    cases = [
        ("pushframe_radial_4096x4096", synth.pushframe_state(n_framelets=256, framelet_height=16, n_samples=4096)),
        ("frame_radial_4096x4096", synth.frame_state(n_lines=4096, n_samples=4096)),
        ("frame_transverse_4096x4096", synth.frame_state(
            n_lines=4096, n_samples=4096, distortion_type=synth.TRANSVERSE,
            dist_coeffs=(0.0, 1.0, 0.0, 1e-6, 2e-6, -1e-6, 1e-8, 0.0, 2e-8, 0.0,
                         0.0, 0.0, 1.0, -1e-6, 1e-6, 2e-6, 0.0, 1e-8, 0.0, -2e-8))),
        ("sar_orbital_64576x2367", synth.sar_state()),
    ]
    for name, state in cases:
        gm = ub.UsgsAstroPlugin(device=dev.index or 0).constructModelFromState(state)
        sz = gm.getImageSize()
        nl, ns = int(sz.line), int(sz.samp)
        n = nl * ns
        rows = torch.arange(nl, device=dev, dtype=torch.float64) + 0.5
        cols = torch.arange(ns, device=dev, dtype=torch.float64) + 0.5
        h = (500.0 * torch.sin(2 * np.pi * rows / 4096.0)[:, None] *
             torch.cos(2 * np.pi * cols / 1024.0)[None, :]).reshape(-1).contiguous()
        res = {}

        def timed(fn):
            fn()
            fn()
            torch.cuda.synchronize()
            ms = []
            for _ in range(5):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                r = fn()
                b.record()
                torch.cuda.synchronize()
                ms.append(a.elapsed_time(b))
            return r, sorted(ms)[len(ms) // 2]

        o1 = tuple(torch.empty(n, dtype=torch.float64, device=dev) for _ in range(3)) + (
            None, torch.empty(n, dtype=torch.int32, device=dev))
        o2 = tuple(torch.empty(n, dtype=torch.float64, device=dev) for _ in range(2)) + (
            None, torch.empty(n, dtype=torch.int32, device=dev))
        (x, y, z, _, st), ms = timed(lambda: gm.imageToGroundGrid(0.5, 0.5, 1.0, 1.0, nl, ns, h, PREC, out=o1))
        res["image_to_ground_mpts"] = n / ms / 1e3
        (l, s, _, st2), ms2 = timed(lambda: gm.groundToImageBatch(x, y, z, PREC, achieved=False, out=o2))
        res["ground_to_image_mpts"] = n / ms2 / 1e3
        res["points"] = n
        res["round_trip_max_line_px"] = float((l.view(nl, ns) - rows[:, None]).abs().max())
        res["round_trip_max_samp_px"] = float((s.view(nl, ns) - cols[None, :]).abs().max())
        res["points_with_flags"] = int(((st & ~_capi.ST_EVALS_MASK) != 0).sum()) + int(((st2 & ~_capi.ST_EVALS_MASK) != 0).sum())
        res["evaluations_per_point"] = float(((st2 >> _capi.ST_EVALS_SHIFT) & 0xff).double().mean())
        out[name] = res
        del gm, x, y, z, l, s, st, st2, h, o1, o2
        torch.cuda.empty_cache()
    return out


def e2e(args, models, x, y, z, seg, n_job, barrier, max_over_ranks):
    """Same metric through the host-pointer C-ABI call: pinned host arrays in, pinned host
    arrays out, host<->device copies inside the timed region (one call per image band)."""
    import torch
    from usgscsm_b200 import _capi
    n = seg * len(models)
    hx, hy, hz = (t.cpu().pin_memory() for t in (x, y, z))
    hl = torch.empty(n, dtype=torch.float64).pin_memory()
    hs = torch.empty(n, dtype=torch.float64).pin_memory()
    hst = torch.empty(n, dtype=torch.int32).pin_memory()

    def call():
        for i, m in enumerate(models):
            o = i * seg
            _capi.check(_capi.lib.usgscsm_b200_ground_to_image(
                m._h, seg, hx.data_ptr() + 8 * o, hy.data_ptr() + 8 * o, hz.data_ptr() + 8 * o, PREC,
                hl.data_ptr() + 8 * o, hs.data_ptr() + 8 * o, None, hst.data_ptr() + 4 * o))

    for _ in range(max(1, min(args.warmup, 2))):
        call()
    barrier()
    t = time.perf_counter()
    for _ in range(args.steps):
        call()
    torch.cuda.synchronize()
    dt = max_over_ranks(time.perf_counter() - t)
    barrier()
    return {"value": n_job * args.steps / dt / 1e6, "unit": "Mpoints/s",
            "h2d_bytes_per_step": 24 * n, "d2h_bytes_per_step": 20 * n,
            "ms_per_step": 1e3 * dt / args.steps,
            "api": "usgscsm_b200_ground_to_image (host pointers, pinned)"}


def e2e_grid(args, model, h, lines, line0):
    """Orthorectification end to end: per-pixel heights up from pinned host memory, ground
    points and status down, pixel coordinates generated in the kernel (28 B/pt down, 8 up)."""
    import torch
    from usgscsm_b200 import _capi
    n = lines * N_SAMPLES
    hh = h[:n].cpu().pin_memory()
    hx, hy, hz = (torch.empty(n, dtype=torch.float64).pin_memory() for _ in range(3))
    hst = torch.empty(n, dtype=torch.int32).pin_memory()

    def call():
        _capi.check(_capi.lib.usgscsm_b200_image_to_ground_grid(
            model._h, line0, 0.5, 1.0, 1.0, lines, N_SAMPLES, hh.data_ptr(), 1, PREC,
            hx.data_ptr(), hy.data_ptr(), hz.data_ptr(), None, hst.data_ptr()))

    call()
    steps = max(1, args.steps // 2)
    t = time.perf_counter()
    for _ in range(steps):
        call()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t
    return {"value": n * steps / dt / 1e6, "unit": "Mpoints/s", "h2d_bytes_per_step": 8 * n,
            "d2h_bytes_per_step": 28 * n, "ms_per_step": 1e3 * dt / steps,
            "api": "usgscsm_b200_image_to_ground_grid (host pointers, pinned)"}


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
