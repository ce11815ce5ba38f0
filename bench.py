#!/usr/bin/env python
"""bench.py -- line-scan groundToImage / imageToGround throughput on B200 (BASELINE.json metric).

N = 1: config C2 of BASELINE.json -- a synthetic orbital NAC-style line-scan image (lroLrocNac
distortion), 5064 samples x 52224 lines, dense-grid orthorectification: imageToGround of every
pixel centre at a synthetic-DEM height, then groundToImage of every resulting ground point.  A
"step" is one groundToImage pass over the whole grid (264.5 M points); imageToGround and the other
configs (C1, C3, C4) are measured beside it under "configs".
N > 1: config C5 -- 64 such images (49368 lines each, 250 M points), images outer, row bands
inner: rank r processes band r of every image; no data-path collective; the gather of one
image's outputs is timed as its own column.

  python bench.py [--gpus N] [--steps K] [--warmup W]          # this engine
  python bench.py --impl reference [...]                        # the reference's CPU code
"""
import argparse
import importlib.util
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
N_LINES_C2 = 52224
C5_IMAGES = 64
C5_LINES = 49368          # 5064 x 49368 = 250.0 M points per image
PREC = 0.001
# algorithmic HBM bytes per point: x,y,z in; line,samp + status out
G2I_BYTES_PER_POINT = 24 + 16 + 4
I2G_GRID_BYTES_PER_POINT = 8 + 24 + 4     # dense grid: height in; x,y,z + status out
CPU_SAMPLE_SECONDS = 4.0


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--lines", type=int, default=0,
                    help="lines per image (default: 52224 at N = 1 = C2, 49368 at N > 1 = C5)")
    ap.add_argument("--images", type=int, default=C5_IMAGES, help="images of the N > 1 job (C5: 64)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-others", action="store_true",
                    help="skip the side measurements (callers, configs C1 / C3 / C4)")
    a = ap.parse_args()
    if a.lines <= 0:
        a.lines = N_LINES_C2 if a.gpus == 1 else C5_LINES
    return a


def synth_module():
    """usgscsm_b200/synth.py loaded by path: the reference arm must not import the package,
    whose __init__ loads the product's shared library."""
    if "usgscsm_b200" in sys.modules:
        from usgscsm_b200 import synth
        return synth
    spec = importlib.util.spec_from_file_location("_ub200_synth", os.path.join(ROOT, "usgscsm_b200", "synth.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def opcounts():
    """Exact FP64 operation counts of the reference per point (oracle/opcount.py ->
    profiles/r2_opcount.json; counting convention of SURVEY.md 8d)."""
    with open(os.path.join(ROOT, "profiles", "r2_opcount.json")) as f:
        return json.load(f)["configs"]


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
    """Image `image` of the job: a C2-shaped line-scan state (5064 x `lines`) on a circular polar
    orbit at 50 km, its own orbit phase and attitude wobble (seed 20261104 + image)."""
    return synth_module().linescan_state(n_lines=lines, n_samples=N_SAMPLES, seed=20261104 + image,
                                         polar=True)


def workload_config(args):
    if args.gpus == 1:
        return {"workload": "C2 orbital line-scan (lroLrocNac distortion) %dx%d dense grid, "
                            "imageToGround at synthetic-DEM heights then groundToImage; value = "
                            "groundToImage" % (N_SAMPLES, args.lines),
                "points_per_step": N_SAMPLES * args.lines, "desired_precision": PREC,
                "sharding": "single GPU",
                "l2": "inputs (6.3 GB) larger than L2; no flush needed"}
    return {"workload": "C5 multi-image batch: %d orbital line-scan images (C2 camera, seeds 1..%d) x "
                        "%dx%d points, images outer / row bands inner; value = groundToImage over "
                        "all images" % (args.images, args.images, N_SAMPLES, args.lines),
            "points_per_step": args.images * N_SAMPLES * args.lines, "desired_precision": PREC,
            "sharding": "rank r takes row band r of every image; no data-path collective; the gather "
                        "of an image's outputs is timed separately (gather)",
            "l2": "inputs per launch larger than L2; no flush needed"}


def peak_hbm():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# ---- the reference's CPU code on a bounded sample (both arms use this one function) ------------
def cpu_sample_rows(n_lines, rows_wanted):
    """A contiguous block of image rows from the middle of the image."""
    rows_wanted = int(max(1, min(n_lines, rows_wanted)))
    first = (n_lines - rows_wanted) // 2
    return first, rows_wanted


def cpu_reference(states, n_lines, direction="g2i", seconds=CPU_SAMPLE_SECONDS, cores=None, proj=None):
    """Reference CPU code (oracle/_ref, OpenMP over points) on a bounded sample of the workload:
    a block of whole image rows from the middle of each given image, sized for about `seconds` of
    work in total.  Returns (again, cores, kind, sample description); again() runs one pass and
    returns (seconds, points)."""
    import numpy as np
    from oracle import ref
    synth = synth_module()
    if cores is None:  # every host core this process may run on -- not OMP_NUM_THREADS (torchrun sets 1)
        cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    rms = [ref.RefModel(s) for s in states]
    if proj is not None:
        for rm, p in zip(rms, proj):
            if p is not None:
                rm.ls_set_projective(True, p)

    def block(first, rows):
        line = np.repeat(np.arange(first, first + rows, dtype=np.float64) + 0.5, N_SAMPLES)
        samp = np.tile(np.arange(N_SAMPLES, dtype=np.float64) + 0.5, rows)
        return line, samp, synth.synthetic_dem(line, samp)

    def prepare(rows):
        work = []
        for rm in rms:
            line, samp, h = block(*cpu_sample_rows(n_lines, rows))
            x, y, z, _, _ = rm.image_to_ground(line, samp, h, PREC, threads=cores)
            work.append((rm, line, samp, h, x, y, z))
        return work

    def one_pass(work):
        t = time.perf_counter()
        for rm, line, samp, h, x, y, z in work:
            if direction == "i2g":
                rm.image_to_ground(line, samp, h, PREC, threads=cores)
            else:
                rm.ground_to_image(x, y, z, PREC, threads=cores)
        return time.perf_counter() - t, sum(w[1].size for w in work)

    probe = prepare(16)
    one_pass(probe)             # threads started, tables in cache
    dt, n = one_pass(probe)
    rows = max(16, int(seconds * n / dt / N_SAMPLES / len(rms)))
    work = prepare(rows)
    first, rows = cpu_sample_rows(n_lines, rows)
    desc = ("rows %d..%d (all %d samples) of %s: %d points per pass, %s, OpenMP over points"
            % (first, first + rows - 1, N_SAMPLES,
               "the image" if len(rms) == 1 else "each of %d images" % len(rms),
               rows * N_SAMPLES * len(rms), "groundToImage" if direction == "g2i" else "imageToGround"))
    return (lambda: one_pass(work)), cores, "reference", desc


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import ref
    if not ref.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libusgscsm_ref.so not built"}))
        return
    if args.gpus == 1:
        states = [workload_state(0, args.lines)]
    else:  # the C5 job: images spread over the batch are sampled (equal shapes, equal cost per point)
        picks = sorted(set(int(round(k * (args.images - 1) / 3.0)) for k in range(4)))
        states = [workload_state(1 + i, args.lines) for i in picks]
    again, cores, kind, desc = cpu_reference(states, args.lines)
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
        "ms_per_step": 1e3 * t_total / args.steps, "higher_is_better": True,
        "scaling": "weak" if args.gpus == 1 else "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args),
        "cpu_baseline": {"value": value, "unit": "Mpoints/s", "cores": cores, "kind": kind,
                         "sample": desc + " (one pass per step)"},
        "e2e": {"value": value, "unit": "Mpoints/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ---- measured denominators ------------------------------------------------------------------
def measure_fp64(device):
    """FP64 pipe of this GPU, measured now by the library's own microbenchmark
    (usgscsm_b200_measure_fp64; profiles/r2_fp64_peak.json holds a longer run): dependent-free DFMA
    stream -> FMA peak; DADD / DMUL alternating -> the issue rate bit-exact (FMA-free) arithmetic
    can reach."""
    import ctypes as C
    from usgscsm_b200 import _capi
    out = {}
    for name, op in (("dfma", 0), ("dadd_dmul_mix", 3)):
        ips, mhz = C.c_double(), C.c_double()
        _capi.check(_capi.lib.usgscsm_b200_measure_fp64(device, op, 400.0, C.byref(ips), C.byref(mhz)))
        out[name] = ips.value
    return {"fma_peak_tflops": 2.0 * out["dfma"] / 1e12, "issue_peak_tinstr": out["dadd_dmul_mix"] / 1e12,
            "how": "usgscsm_b200_measure_fp64, 0.4 s per instruction kind, this run"}


def measure_link(dev, barrier, nbytes=1 << 30, reps=4):
    """Pinned host <-> device copy rates of this rank's GPU, each direction alone and both at
    once (what the host-pointer pipeline does): the ceiling of every e2e number."""
    import torch
    h_in = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    d_in = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    d_out = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def run(up, down):
        barrier()
        t = time.perf_counter()
        for _ in range(reps):
            if up:
                with torch.cuda.stream(s1):
                    d_in.copy_(h_in, non_blocking=True)
            if down:
                with torch.cuda.stream(s2):
                    h_out.copy_(d_out, non_blocking=True)
        s1.synchronize()
        s2.synchronize()
        return nbytes * reps / (time.perf_counter() - t) / 1e9

    run(True, True)
    res = {"h2d_gbs_alone": run(True, False), "d2h_gbs_alone": run(False, True)}
    both = run(True, True)
    res["h2d_gbs_duplex"] = res["d2h_gbs_duplex"] = both
    # the DEM-grid call's own traffic mix, copies only: 4 bytes up for every 20 down, both at once
    up_view, up_dst = h_in[:nbytes // 5], d_in[:nbytes // 5]
    barrier()
    t = time.perf_counter()
    for _ in range(reps):
        with torch.cuda.stream(s1):
            up_dst.copy_(up_view, non_blocking=True)
        with torch.cuda.stream(s2):
            h_out.copy_(d_out, non_blocking=True)
    s1.synchronize()
    s2.synchronize()
    res["d2h_gbs_mix_4_up_20_down"] = nbytes * reps / (time.perf_counter() - t) / 1e9
    # The same mix for a fixed WINDOW of time instead of a fixed volume: with several ranks on one
    # host the fixed-volume figures above flatter the slower links (the faster ranks finish first
    # and leave them the host to themselves); here every rank copies until the window closes.
    piece = min(nbytes, 256 << 20)
    window = 0.75
    barrier()
    t = time.perf_counter()
    moved = 0
    while True:
        with torch.cuda.stream(s1):
            d_in[:piece // 5].copy_(h_in[:piece // 5], non_blocking=True)
        with torch.cuda.stream(s2):
            h_out[:piece].copy_(d_out[:piece], non_blocking=True)
        s2.synchronize()
        s1.synchronize()
        moved += piece
        dt = time.perf_counter() - t
        if dt >= window:
            break
    res["d2h_gbs_mix_sustained"] = moved / dt / 1e9
    del h_in, h_out, d_in, d_out
    return res


def ncu_summary():
    """The committed ncu --set full capture of the dominant kernel (profiles/ncu_g2i.json)."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_g2i.json")) as f:
            return json.load(f)
    except (OSError, ValueError):
        return None


def timed_launches(fn, steps, barrier, max_over_ranks):
    """K calls of fn bracketed by barrier + synchronize; CUDA events on the launching stream.
    Returns (total ms, max over ranks; mean ms per call on this rank)."""
    import torch
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
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


def median_ms(fn, reps=5, warm=2):
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    return sorted(ms)[len(ms) // 2]


def footprint_grid(model, n_rows, n_cols, dev, shrink=0.01):
    """GeoGrid of a north-up longlat raster of n_rows x n_cols pixels lying inside the image
    footprint of `model` (polar orbit: the footprint is a lon / lat rectangle up to the swath's
    curvature; the raster is the largest such rectangle inside it, shrunk by 1 %)."""
    import numpy as np
    import torch
    import usgscsm_b200 as ub
    sz = model.getImageSize()
    nl, ns = float(sz.line), float(sz.samp)
    # corners and edge midpoints: (line, samp)
    cl = torch.tensor([0.0, 0.0, nl, nl, nl / 2, nl / 2, 0.0, nl], dtype=torch.float64, device=dev)
    cs = torch.tensor([0.0, ns, 0.0, ns, 0.0, ns, ns / 2, ns / 2], dtype=torch.float64, device=dev)
    x, y, z, _, _ = model.imageToGroundBatch(cl, cs, 0.0)
    x, y, z = x.cpu().numpy(), y.cpu().numpy(), z.cpu().numpy()
    lon = np.degrees(np.arctan2(y, x))
    lat = np.degrees(np.arcsin(z / np.sqrt(x * x + y * y + z * z)))

    def inner(v, edge_a, edge_b):
        a, b = v[edge_a], v[edge_b]
        if np.median(a) > np.median(b):
            a, b = b, a
        return a.max(), b.min()
    west, east = inner(lon, [0, 2, 4], [1, 3, 5])      # the samp = 0 and samp = ns edges
    south, north = inner(lat, [0, 1, 6], [2, 3, 7])    # the line = 0 and line = nl edges
    dlo, dla = (east - west) * shrink, (north - south) * shrink
    west, east, south, north = west + dlo, east - dlo, south + dla, north - dla
    gt = [west, (east - west) / n_cols, 0.0, north, 0.0, -(north - south) / n_rows]
    a, b = model.getEllipsoid()
    return ub.GeoGrid.fromProjString("+proj=longlat +a=%.17g +b=%.17g +no_defs" % (a, b), gt)


def footprint_dem(model, n_rows, n_cols, dev):
    """A float32 DEM raster (the synthetic DEM) on footprint_grid, on the device."""
    import numpy as np
    import torch
    import usgscsm_b200 as ub
    grid = footprint_grid(model, n_rows, n_cols, dev)
    h = torch.empty((n_rows, n_cols), dtype=torch.float32, device=dev)
    cols = torch.arange(n_cols, device=dev, dtype=torch.float64) + 0.5
    cterm = torch.cos(2 * np.pi * cols / 1024.0)
    for r0 in range(0, n_rows, 4096):
        r1 = min(n_rows, r0 + 4096)
        rows = torch.arange(r0, r1, device=dev, dtype=torch.float64) + 0.5
        h[r0:r1] = (500.0 * torch.sin(2 * np.pi * rows / 4096.0)[:, None] * cterm[None, :]).float()
    return ub.Dem(h, grid)


def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import usgscsm_b200 as ub
    from usgscsm_b200 import _capi, sharding

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the engine has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    host = {}
    if world > 1:
        if os.environ.get("USGSCSM_B200_NUMA_BIND", "1") != "0":
            host["numa_bind"] = sharding.bind_to_gpu_numa_node(local)
        # stdout carries ONE JSON line: keep NCCL's version banner (NCCL_DEBUG=VERSION) off it
        if os.environ.get("NCCL_DEBUG", "").upper() not in ("INFO", "TRACE"):
            os.environ.pop("NCCL_DEBUG", None)
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

    def sum_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    def gather_floats(v):
        """[v of rank 0, v of rank 1, ...] on every rank"""
        if world == 1:
            return [float(v)]
        t = torch.zeros(world, dtype=torch.float64, device=dev)
        t[rank] = v
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return [float(x) for x in t.tolist()]

    ctx = dict(args=args, rank=rank, world=world, local=local, dev=dev, barrier=barrier,
               max_over_ranks=max_over_ranks, sum_over_ranks=sum_over_ranks, gather_floats=gather_floats,
               host=host)
    result = run_single(ctx) if world == 1 else run_c5(ctx)
    if rank == 0:
        print(json.dumps(result))
    if world > 1:
        dist.destroy_process_group()


def roofline_object(kern_ms, n, flops_per_point, flops_note, fp64, clocks):
    """FP64 roofline of the dominant kernel (lsGroundToImageKernel): algorithmic flops = the
    reference's own executed FP64 operations per point (op-counted) x points / kernel time."""
    peak_hbm_gbs, peak_src = peak_hbm()
    ncu = ncu_summary()
    achieved = flops_per_point * n / (kern_ms * 1e-3) / 1e12
    hbm = G2I_BYTES_PER_POINT * n / (kern_ms * 1e-3) / 1e9
    traffic = None
    executed = None
    if ncu:
        try:
            traffic = (ncu["dram_bytes_read"] + ncu["dram_bytes_write"]) / ncu["points"] * n / 1e9
            executed = {"fp64_pipe_active_pct": ncu.get("fp64_pipe_active_pct"),
                        "issue_frac_ncu": ncu["fp64_thread_inst_per_cycle"] / ncu["fp64_issue_peak_per_cycle"],
                        "source": "profiles/ncu_g2i.json (%s)" % ncu.get("capture", "ncu --set full")}
        except KeyError:
            pass
    return {
        "bound": "fp64", "achieved": achieved, "peak": fp64["fma_peak_tflops"], "unit": "TFLOP/s",
        "frac": achieved / fp64["fma_peak_tflops"],
        "traffic": traffic, "traffic_unit": "GB of DRAM traffic per launch (ncu dram__bytes_read + write)",
        "kernel": "lsGroundToImageKernel", "kernel_ms": kern_ms, "points_per_launch": n,
        "flops_per_point": flops_per_point, "flops_source": flops_note,
        "peak_source": "measured: " + fp64["how"] + " (DFMA thread-instructions/s x 2)",
        "no_fma": {"peak": fp64["issue_peak_tinstr"], "unit": "T FP64 thread-instructions/s",
                   "frac": achieved / fp64["issue_peak_tinstr"],
                   "note": "the reference's arithmetic has no FMA contraction; kernels that round like it "
                           "issue one instruction per flop, so the measured DADD/DMUL issue rate is the "
                           "ceiling for executed work.  frac > what the pipe could issue means the kernel "
                           "executes fewer operations than the reference does for the same result "
                           "(hoisted reciprocals, shared Lagrange weights)."},
        "executed": executed,
        "hbm": {"bound": "hbm", "achieved": hbm, "peak": peak_hbm_gbs, "unit": "GB/s", "frac": hbm / peak_hbm_gbs,
                "bytes_per_point": G2I_BYTES_PER_POINT, "peak_source": peak_src,
                "note": "not the binding roofline: arithmetic intensity is ~40 flop/B"},
        "clock_mhz_under_load": (clocks or {}).get("sm_mhz"),
    }


def run_single(ctx):
    """N = 1: config C2 on one GPU, the other configs beside it."""
    import numpy as np
    import torch
    import usgscsm_b200 as ub
    from usgscsm_b200 import _capi
    args, dev, local = ctx["args"], ctx["dev"], ctx["local"]
    barrier, max_over_ranks = ctx["barrier"], ctx["max_over_ranks"]
    lines = args.lines
    n = lines * N_SAMPLES
    state = workload_state(0, lines)
    plugin = ub.UsgsAstroPlugin(device=local)
    model = plugin.constructModelFromState(state)
    stream = torch.cuda.current_stream(dev).cuda_stream
    lib, hm = _capi.lib, model._h

    # synthetic DEM heights of the pixel grid
    h = torch.empty(n, dtype=torch.float64, device=dev)
    cols = torch.arange(N_SAMPLES, device=dev, dtype=torch.float64) + 0.5
    cterm = torch.cos(2 * np.pi * cols / 1024.0)
    for r0 in range(0, lines, 4096):
        r1 = min(lines, r0 + 4096)
        rows = torch.arange(r0, r1, device=dev, dtype=torch.float64) + 0.5
        h[r0 * N_SAMPLES:r1 * N_SAMPLES] = (500.0 * torch.sin(2 * np.pi * rows / 4096.0)[:, None] * cterm[None, :]).reshape(-1)
    x, y, z = (torch.empty(n, dtype=torch.float64, device=dev) for _ in range(3))
    st_g = torch.empty(n, dtype=torch.int32, device=dev)
    out_line, out_samp = (torch.empty(n, dtype=torch.float64, device=dev) for _ in range(2))
    st_i = torch.empty(n, dtype=torch.int32, device=dev)
    p = lambda t: t.data_ptr()

    def i2g():
        _capi.check(lib.usgscsm_b200_image_to_ground_grid_dev(
            hm, 0.5, 0.5, 1.0, 1.0, lines, N_SAMPLES, p(h), 1, PREC, p(x), p(y), p(z), None, p(st_g), stream))

    def g2i(prec=PREC):
        _capi.check(lib.usgscsm_b200_ground_to_image_dev(
            hm, n, p(x), p(y), p(z), prec, p(out_line), p(out_samp), None, p(st_i), stream))

    fp64 = measure_fp64(local)
    i2g()
    torch.cuda.synchronize()
    for _ in range(args.warmup):
        g2i()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = lib.usgscsm_b200_kernel_launches()
    total_ms, kern_ms = timed_launches(g2i, args.steps, barrier, max_over_ranks)
    launches = lib.usgscsm_b200_kernel_launches() - launches0
    clocks = sampler.stop()
    value = n * args.steps / (total_ms * 1e-3) / 1e6

    # correctness of what was just timed: the round trip must close and no flag may be set
    rows_all = (torch.arange(lines, device=dev, dtype=torch.float64) + 0.5)[:, None]
    err = max(float((out_line.view(lines, N_SAMPLES) - rows_all).abs().max()),
              float((out_samp.view(lines, N_SAMPLES) - cols[None, :]).abs().max()))
    obs = _capi.ST_REFERENCE_MASK
    flags_obs = int(((st_i & obs) != 0).sum().item()) + int(((st_g & obs) != 0).sum().item())
    flags_diag = int(((st_i & ~obs & ~_capi.ST_EVALS_MASK) != 0).sum().item())
    mean_evals = float(((st_i >> _capi.ST_EVALS_SHIFT) & 0xff).double().mean().item())

    ops = opcounts()
    lin = ops["C2_linescan_nac"]["ground_to_image_model"]
    flops_pt = lin["flops_per_evaluation"] * mean_evals + lin["flops_fixed"]
    result = {
        "metric": METRIC, "value": value, "unit": "Mpoints/s", "n_gpus": 1,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": workload_config(args), "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofline_object(
            kern_ms, n, flops_pt,
            "profiles/r2_opcount.json (ptrace op count of the unmodified reference binary): %.1f per "
            "detector-view evaluation x %.3f evaluations/pt measured by the kernel + %.1f fixed"
            % (lin["flops_per_evaluation"], mean_evals, lin["flops_fixed"]), fp64, clocks),
        "check": {"round_trip_max_px": err, "points_with_reference_flags": flags_obs,
                  "points_with_diagnostic_flags_only": flags_diag,
                  "evaluations_per_point": mean_evals},
    }

    # imageToGround beside it (dense grid: row / column tables + per-pixel kernel)
    i2g_total_ms, i2g_kern_ms = timed_launches(i2g, args.steps, barrier, max_over_ranks)
    i2g_flops = ops["C2_linescan_nac"]["image_to_ground"]["flops_per_point"]
    result["image_to_ground"] = {
        "value": n * args.steps / (i2g_total_ms * 1e-3) / 1e6, "unit": "Mpoints/s",
        "ms_per_step": i2g_total_ms / args.steps, "kernel_ms": i2g_kern_ms,
        "roofline": {"bound": "fp64", "flops_per_point": i2g_flops,
                     "achieved": i2g_flops * n / (i2g_kern_ms * 1e-3) / 1e12, "peak": fp64["fma_peak_tflops"],
                     "unit": "TFLOP/s", "frac": i2g_flops * n / (i2g_kern_ms * 1e-3) / 1e12 / fp64["fma_peak_tflops"],
                     "note": "flops = what the reference executes per pixel; the grid kernel shares the "
                             "per-row and per-column parts, so it executes far fewer",
                     "hbm_gbs": I2G_GRID_BYTES_PER_POINT * n / (i2g_kern_ms * 1e-3) / 1e9,
                     "hbm_frac": I2G_GRID_BYTES_PER_POINT * n / (i2g_kern_ms * 1e-3) / 1e9 / peak_hbm()[0]}}

    # second column: desiredPrecision 1e-8, what the reference's own CLI test uses
    g2i(1e-8)
    tight_total_ms, tight_kern_ms = timed_launches(lambda: g2i(1e-8), args.steps, barrier, max_over_ranks)
    ev8 = float(((st_i >> _capi.ST_EVALS_SHIFT) & 0xff).double().mean().item())
    f8 = lin["flops_per_evaluation"] * ev8 + lin["flops_fixed"]
    result["ground_to_image_prec_1e-8"] = {
        "value": n * args.steps / (tight_total_ms * 1e-3) / 1e6, "unit": "Mpoints/s", "kernel_ms": tight_kern_ms,
        "evaluations_per_point": ev8, "flops_per_point": f8,
        "fp64_frac": f8 * n / (tight_kern_ms * 1e-3) / 1e12 / fp64["fma_peak_tflops"]}

    # imageToGround from explicit (line, samp) arrays (not the dense-grid tables): 52 B/pt
    ex_line = (torch.arange(lines, device=dev, dtype=torch.float64) + 0.5).repeat_interleave(N_SAMPLES)
    ex_samp = cols.repeat(lines)
    ex_ms = median_ms(lambda: _capi.check(lib.usgscsm_b200_image_to_ground_dev(
        hm, n, p(ex_line), p(ex_samp), p(h), 1, PREC, p(out_line), p(out_samp), p(ex_line), None,
        p(st_i), stream)), reps=3, warm=1)
    result["image_to_ground_explicit"] = {"value": n / ex_ms / 1e3, "unit": "Mpoints/s", "kernel_ms": ex_ms}
    del ex_line, ex_samp

    # orthorectification as callers run it: groundToImage over a DEM raster, ground points
    # generated in the kernel (float32 heights in, line / samp / status out)
    dem = footprint_dem(model, lines, N_SAMPLES, dev)
    dg_ms = median_ms(lambda: _capi.check(lib.usgscsm_b200_ground_to_image_dem_grid_dev(
        hm, dem.c, 0, 0, lines, N_SAMPLES, PREC, p(out_line), p(out_samp), None, p(st_i), stream)),
        reps=max(3, args.steps), warm=2)
    inside = float(((st_i & _capi.ST_OUT_OF_IMAGE) == 0).double().mean().item())
    result["ground_to_image_dem_grid"] = {
        "value": n / dg_ms / 1e3, "unit": "Mpoints/s", "kernel_ms": dg_ms,
        "hbm_bytes_per_point": 4 + 16 + 4, "fraction_inside_image": inside,
        "evaluations_per_point": float(((st_i >> _capi.ST_EVALS_SHIFT) & 0xff).double().mean().item()),
        "api": "usgscsm_b200_ground_to_image_dem_grid_dev (longlat DEM %dx%d float32 inside the footprint)"
               % (lines, N_SAMPLES)}

    if not args.no_others:
        result["callers"] = callers_side(model, x, y, z, min(n, 1 << 26), dev, lines)

    if not args.no_e2e:
        link = measure_link(dev, barrier)
        result["e2e"] = e2e_dem_grid(ctx, [model], [dem], lines, n, link)
        result["e2e_explicit_arrays"] = e2e_explicit(args, model, x, y, z, n, link)
        result["e2e_image_to_ground_grid"] = e2e_i2g_grid(args, model, h, lines, link)
        result["e2e_pageable_numpy"] = e2e_pageable(model, x, y, z, min(n, 1 << 26))
    del dem

    if not args.no_cpu:
        from oracle import ref
        if ref.available():
            proj = [model.lsDerived()[5:19]]
            again, cores, kind, desc = cpu_reference([state], lines, proj=proj)
            again()
            t_total, n_total = 0.0, 0
            for _ in range(2):
                dt, k = again()
                t_total += dt
                n_total += k
            result["cpu_baseline"] = {"value": n_total / t_total / 1e6, "unit": "Mpoints/s", "cores": cores,
                                      "kind": kind, "sample": desc + " (2 passes after 1 warm-up)"}
            again1, _, _, desc1 = cpu_reference([state], lines, seconds=2.0, cores=1, proj=proj)
            dt1, k1 = again1()
            result["cpu_baseline"]["single_thread"] = {"value": k1 / dt1 / 1e6, "sample": desc1}
            againi, _, _, desci = cpu_reference([state], lines, direction="i2g", seconds=2.0, proj=proj)
            againi()
            dti, ki = againi()
            result["image_to_ground"]["cpu_baseline"] = {"value": ki / dti / 1e6, "unit": "Mpoints/s",
                                                         "cores": cores, "kind": kind, "sample": desci}
    del x, y, z, h, out_line, out_samp, st_g, st_i
    torch.cuda.empty_cache()
    if not args.no_others:
        result["configs"] = other_configs(dev, fp64, ops, not args.no_cpu)
        # the N > 1 job (C5) on this one GPU, same timing method: the 1-GPU point of its strong scaling
        import copy
        a5 = copy.copy(args)
        a5.lines, a5.no_e2e, a5.steps, a5.warmup = C5_LINES, True, 1, 3
        c5 = run_c5(dict(ctx, args=a5))
        torch.cuda.empty_cache()
        result["configs"]["C5_64_images_on_one_gpu"] = {
            "value": c5["value"], "unit": "Mpoints/s", "ms_per_step": c5["ms_per_step"],
            "evaluations_per_point": c5["check"]["evaluations_per_point"],
            "roofline_frac": c5["roofline"]["frac"], "job": c5["job"],
            "note": "bench.py --gpus N (N > 1) runs this job; its images need 3.0 to 5.2 evaluations per point "
                    "(mean above C2's 3.04) because the reference's projective seed is poorer for some orbits"}
    return result


def callers_side(model, x, y, z, m, dev, lines):
    """Side measurements of the batched caller-level entry points on the C2 line-scan model
    (SURVEY.md 8f N3).  Device resident, median of 3 after 1 warm-up."""
    import torch
    from usgscsm_b200 import _capi
    lib, h = _capi.lib, model._h
    stream = torch.cuda.current_stream(dev).cuda_stream
    f64 = lambda k: [torch.empty(m, dtype=torch.float64, device=dev) for _ in range(k)]
    st = torch.empty(m, dtype=torch.int32, device=dev)
    part = torch.empty(6 * m, dtype=torch.float64, device=dev)
    o = f64(7)
    rows = torch.arange(lines, device=dev, dtype=torch.float64) + 0.5
    line = rows.repeat_interleave(N_SAMPLES)[:m].contiguous()
    samp = (torch.arange(N_SAMPLES, device=dev, dtype=torch.float64) + 0.5).repeat(lines)[:m].contiguous()
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
    out = {"points": m, "unit": "Mpoints/s", "points_all_": m4,
           "searches_per_point": {"ground_partials": 4, "sensor_partials_index0": 2,
                                  "all_sensor_partials_%d_params" % npar: npar + 1,
                                  "all_ground_to_image_covar": npar + 6}}
    for name, fn in calls.items():
        ms = median_ms(lambda: _capi.check(fn()), reps=3, warm=1)
        out[name] = (m4 if name.startswith("all_") else m) / ms / 1e3
    return out


def other_configs(dev, fp64, ops, with_cpu):
    """Configs C1, C3 and C4 of BASELINE.json beside the headline: device-resident rates with a
    per-config roofline (op-counted flops of the reference; the frame model against HBM) and the
    reference's CPU rate on a bounded sample."""
    import numpy as np
    import torch
    import usgscsm_b200 as ub
    from usgscsm_b200 import _capi, synth
    out = {}
    hbm_peak = peak_hbm()[0]
    transverse = (0.0, 1.0, 0.0, 1e-6, 2e-6, -1e-6, 1e-8, 0.0, 2e-8, 0.0,
                  0.0, 0.0, 1.0, -1e-6, 1e-6, 2e-6, 0.0, 1e-8, 0.0, -2e-8)
    cases = [
        ("C3_pushframe_radial_4096x4096", synth.pushframe_state(n_framelets=256, framelet_height=16, n_samples=4096),
         "C3_pushframe_radial", "fp64"),
        ("C3_pushframe_transverse_4096x4096", synth.pushframe_state(
            n_framelets=256, framelet_height=16, n_samples=4096, distortion_type=synth.TRANSVERSE,
            dist_coeffs=transverse), "C3_pushframe_radial", "fp64"),
        ("C3_frame_radial_4096x4096", synth.frame_state(n_lines=4096, n_samples=4096), "C3_frame_radial", "hbm"),
        ("C3_frame_transverse_4096x4096", synth.frame_state(
            n_lines=4096, n_samples=4096, distortion_type=synth.TRANSVERSE, dist_coeffs=transverse),
         "C3_frame_radial", "hbm"),
    ]
    plugin = ub.UsgsAstroPlugin(device=dev.index or 0)
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)

    def cpu_rates(state, line, samp, hh, x, y, z, seconds=1.5):
        """Reference CPU rates on the first points of the arrays, ~`seconds` each direction."""
        from oracle import ref
        if not (with_cpu and ref.available()):
            return None
        rm = ref.RefModel(state)
        k = min(line.size, 20000)
        t = time.perf_counter()
        rm.ground_to_image(x[:k], y[:k], z[:k], PREC, threads=cores)
        rate = k / (time.perf_counter() - t)
        k = int(min(line.size, max(20000, rate * seconds)))
        t = time.perf_counter()
        rm.ground_to_image(x[:k], y[:k], z[:k], PREC, threads=cores)
        g = k / (time.perf_counter() - t) / 1e6
        t = time.perf_counter()
        rm.image_to_ground(line[:k], samp[:k], hh[:k], PREC, threads=cores)
        i = k / (time.perf_counter() - t) / 1e6
        return {"ground_to_image": g, "image_to_ground": i, "unit": "Mpoints/s", "cores": cores,
                "kind": "reference", "sample": "the first %d points of the workload, one pass per direction" % k}

    def roof(kind, flops, bytes_pt, npts, ms):
        if kind == "hbm":
            a = bytes_pt * npts / (ms * 1e-3) / 1e9
            return {"bound": "hbm", "achieved": a, "peak": hbm_peak, "unit": "GB/s", "frac": a / hbm_peak,
                    "bytes_per_point": bytes_pt}
        a = flops * npts / (ms * 1e-3) / 1e12
        return {"bound": "fp64", "achieved": a, "peak": fp64["fma_peak_tflops"], "unit": "TFLOP/s",
                "frac": a / fp64["fma_peak_tflops"], "flops_per_point": flops}

    for name, state, opkey, bound in cases:
        gm = plugin.constructModelFromState(state)
        sz = gm.getImageSize()
        nl, ns = int(sz.line), int(sz.samp)
        n = nl * ns
        rows = torch.arange(nl, device=dev, dtype=torch.float64) + 0.5
        cols = torch.arange(ns, device=dev, dtype=torch.float64) + 0.5
        h = (500.0 * torch.sin(2 * np.pi * rows / 4096.0)[:, None] *
             torch.cos(2 * np.pi * cols / 1024.0)[None, :]).reshape(-1).contiguous()
        o1 = tuple(torch.empty(n, dtype=torch.float64, device=dev) for _ in range(3)) + (
            None, torch.empty(n, dtype=torch.int32, device=dev))
        o2 = tuple(torch.empty(n, dtype=torch.float64, device=dev) for _ in range(2)) + (
            None, torch.empty(n, dtype=torch.int32, device=dev))
        ms1 = median_ms(lambda: gm.imageToGroundGrid(0.5, 0.5, 1.0, 1.0, nl, ns, h, PREC, out=o1))
        x, y, z, _, st = o1
        ms2 = median_ms(lambda: gm.groundToImageBatch(x, y, z, PREC, achieved=False, out=o2))
        l, s, _, st2 = o2
        obs = _capi.ST_REFERENCE_MASK
        res = {"points": n,
               "image_to_ground": {"value": n / ms1 / 1e3, "unit": "Mpoints/s", "kernel_ms": ms1,
                                   "roofline": roof(bound, ops[opkey]["image_to_ground"]["flops_per_point"],
                                                    I2G_GRID_BYTES_PER_POINT, n, ms1)},
               "ground_to_image": {"value": n / ms2 / 1e3, "unit": "Mpoints/s", "kernel_ms": ms2,
                                   "roofline": roof(bound, ops[opkey]["ground_to_image_prec_0.001"]["flops_per_point"],
                                                    G2I_BYTES_PER_POINT, n, ms2)},
               "round_trip_max_line_px": float((l.view(nl, ns) - rows[:, None]).abs().max()),
               "round_trip_max_samp_px": float((s.view(nl, ns) - cols[None, :]).abs().max()),
               "points_with_reference_flags": int(((st & obs) != 0).sum()) + int(((st2 & obs) != 0).sum()),
               "evaluations_per_point": float(((st2 >> _capi.ST_EVALS_SHIFT) & 0xff).double().mean())}
        if "transverse" in name:
            res["note"] = "flops per point op-counted on the radial variant of the same model"
        else:
            # 16 images of this size in ONE call (usgscsm_b200_*_images_dev: kernels forked over side
            # streams, one image's tail under the next one's body)
            K = 16
            maker = (lambda k: synth.pushframe_state(n_framelets=256, framelet_height=16, n_samples=4096, seed=3 + k)) \
                if "pushframe" in name else (lambda k: synth.frame_state(n_lines=4096, n_samples=4096, seed=5 + k))
            many = [gm] + [plugin.constructModelFromState(maker(k)) for k in range(1, K)]
            hk = h.repeat(K)
            m1 = tuple(torch.empty(n * K, dtype=torch.float64, device=dev) for _ in range(3)) + (
                None, torch.empty(n * K, dtype=torch.int32, device=dev))
            m2 = tuple(torch.empty(n * K, dtype=torch.float64, device=dev) for _ in range(2)) + (
                None, torch.empty(n * K, dtype=torch.int32, device=dev))
            msi = median_ms(lambda: ub.imageToGroundGridImages(many, 0.5, 0.5, 1.0, 1.0, [nl] * K, [ns] * K, hk, PREC, out=m1))
            msg = median_ms(lambda: ub.groundToImageImages(many, [n] * K, m1[0], m1[1], m1[2], PREC, out=m2))
            res["sixteen_images_per_call"] = {
                "points": n * K,
                "image_to_ground": {"value": n * K / msi / 1e3, "unit": "Mpoints/s", "ms": msi,
                                    "hbm_frac": I2G_GRID_BYTES_PER_POINT * n * K / (msi * 1e-3) / 1e9 / hbm_peak,
                                    "fp64_frac": ops[opkey]["image_to_ground"]["flops_per_point"] * n * K / (msi * 1e-3) / 1e12 / fp64["fma_peak_tflops"]},
                "ground_to_image": {"value": n * K / msg / 1e3, "unit": "Mpoints/s", "ms": msg,
                                    "hbm_frac": G2I_BYTES_PER_POINT * n * K / (msg * 1e-3) / 1e9 / hbm_peak,
                                    "fp64_frac": ops[opkey]["ground_to_image_prec_0.001"]["flops_per_point"] * n * K / (msg * 1e-3) / 1e12 / fp64["fma_peak_tflops"]},
                "api": "usgscsm_b200_image_to_ground_grid_images_dev / usgscsm_b200_ground_to_image_images_dev"}
            del many, hk, m1, m2
        k = 400000
        idx = torch.arange(0, n, max(1, n // k), device=dev)[:k]
        cpu = cpu_rates(state, (idx // ns).double().cpu().numpy() + 0.5, (idx % ns).double().cpu().numpy() + 0.5,
                        h[idx].cpu().numpy(), x[idx].cpu().numpy(), y[idx].cpu().numpy(), z[idx].cpu().numpy())
        if cpu:
            res["cpu_baseline"] = cpu
        out[name] = res
        del gm, x, y, z, l, s, st, st2, h, o1, o2
        torch.cuda.empty_cache()

    # C4: SAR, 100 M unordered DEM-sampled ground points, 0.1 % outside the Doppler bracket
    state = synth.sar_state()
    gm = plugin.constructModelFromState(state)
    sz = gm.getImageSize()
    nl, ns = int(sz.line), int(sz.samp)
    n = 100_000_000
    gen = np.random.Generator(np.random.MT19937(4))   # mt19937, seed 4
    line = torch.as_tensor(gen.uniform(0.0, nl, n), device=dev)
    samp = torch.as_tensor(gen.uniform(0.0, ns, n), device=dev)
    hh = torch.as_tensor(gen.uniform(-1000.0, 1000.0, n), device=dev)
    x, y, z, _, st = gm.imageToGroundBatch(line, samp, hh, PREC, achieved=False)
    # 0.1 % of the points moved far along track, outside the image's time span: no Doppler bracket
    bad = torch.as_tensor(gen.choice(n, n // 1000, replace=False), device=dev)
    far = gm.imageToGroundBatch(line[bad] + 3.0 * nl, samp[bad], hh[bad], PREC, achieved=False)
    x[bad], y[bad], z[bad] = far[0], far[1], far[2]
    o2 = tuple(torch.empty(n, dtype=torch.float64, device=dev) for _ in range(2)) + (
        None, torch.empty(n, dtype=torch.int32, device=dev))
    ms_i = median_ms(lambda: gm.imageToGroundBatch(line, samp, hh, PREC, achieved=False), reps=3, warm=1)
    ms_g = median_ms(lambda: gm.groundToImageBatch(x, y, z, PREC, achieved=False, out=o2), reps=3, warm=1)
    l, s, _, st2 = o2
    good = torch.ones(n, dtype=torch.bool, device=dev)
    good[bad] = False
    failed = (st2 & (_capi.ST_ERROR_UNKNOWN | _capi.ST_ERROR_ALGORITHM | _capi.ST_ERROR_OTHER)) != 0
    fl = ops["C4_sar_orbital"]
    res = {"points": n, "input": "i2g of uniformly random (line, samp) in the image (mt19937 seed 4) at heights "
                                 "uniform in [-1000, 1000] m, unordered; 0.1 % replaced by points 3 image "
                                 "lengths along track (no Doppler bracket)",
           "ground_to_image": {"value": n / ms_g / 1e3, "unit": "Mpoints/s", "kernel_ms": ms_g,
                               "roofline": roof("fp64", fl["ground_to_image_prec_0.001"]["flops_per_point"],
                                                G2I_BYTES_PER_POINT, n, ms_g)},
           "image_to_ground": {"value": n / ms_i / 1e3, "unit": "Mpoints/s", "kernel_ms": ms_i,
                               "roofline": roof("fp64", fl["image_to_ground"]["flops_per_point"], 52, n, ms_i)},
           "round_trip_max_line_px": float((l[good] - line[good]).abs().max()),
           "round_trip_max_samp_px": float((s[good] - samp[good]).abs().max()),
           "points_failing_as_planned": int(failed[bad].sum()), "points_planned_to_fail": int(bad.numel()),
           "other_points_with_error_flags": int(failed[good].sum()),
           "points_with_diagnostic_flags_only": int((((st2 & ~_capi.ST_REFERENCE_MASK & ~_capi.ST_EVALS_MASK) != 0) & ~failed).sum()),
           "diagnostic_note": "ST_ITERATION_CAP from Brent's method is a diagnostic the reference does not report: "
                              "its brentRoot returns the last iterate after 30 steps without a warning "
                              "(src/Utilities.cpp:545-622)",
           "evaluations_per_point": float(((st2 >> _capi.ST_EVALS_SHIFT) & 0xff).double().mean())}
    k = 200000
    cpu = cpu_rates(state, line[:k].cpu().numpy(), samp[:k].cpu().numpy(), hh[:k].cpu().numpy(),
                    x[:k].cpu().numpy(), y[:k].cpu().numpy(), z[:k].cpu().numpy(), seconds=2.0)
    if cpu:
        res["cpu_baseline"] = cpu
    out["C4_sar_100M_random_ground_points"] = res
    del gm, line, samp, hh, x, y, z, l, s, st, st2, o2
    torch.cuda.empty_cache()

    # C1: the reference's constant-velocity line-scan ISD, 16 x 16, image -> ground -> image
    isd = os.path.join(ROOT, "tests", "golden", "ref_data", "constVelocityLineScan.img")
    try:
        gm = plugin.constructModelFromISD(isd, "USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL")
        samp, line = np.meshgrid(np.arange(16.0), np.arange(16.0), indexing="ij")   # bin/usgscsm_cam_test.cc:294-309
        x, y, z, _, st = gm.imageToGroundBatch(line.reshape(-1), samp.reshape(-1), 0.0, PREC)
        l, s, _, st2 = gm.groundToImageBatch(x, y, z, PREC)
        err = np.hypot(l - line.reshape(-1), s - samp.reshape(-1))
        c1 = {"points": 256, "round_trip_px": {"min": float(err.min()), "median": float(np.median(err)),
                                               "max": float(err.max())}}
        from oracle import ref
        if with_cpu and ref.available():
            rm = ref.RefModel.from_isd(isd)
            rx, ry, rz, _, _ = rm.image_to_ground(line.reshape(-1), samp.reshape(-1), np.zeros(256), PREC)
            rl, rs, _, _ = rm.ground_to_image(rx, ry, rz, PREC)
            c1["max_abs_difference_from_reference"] = {
                "image_to_ground_m": float(max(np.abs(rx - x).max(), np.abs(ry - y).max(), np.abs(rz - z).max())),
                "ground_to_image_px": float(max(np.abs(rl - l).max(), np.abs(rs - s).max()))}
        out["C1_constVelocityLineScan_isd_16x16"] = c1
    except Exception as e:   # the fixture travels with the repo; report rather than hide
        out["C1_constVelocityLineScan_isd_16x16"] = {"error": str(e)}
    return out


# ---- end to end through the host-pointer C-ABI ---------------------------------------------------
def link_fraction(h2d_bytes, d2h_bytes, seconds, link):
    """Share of the measured duplex link rate the call sustained, for the direction that binds."""
    up, down = h2d_bytes / seconds / 1e9, d2h_bytes / seconds / 1e9
    res = {"h2d_gbs": up, "d2h_gbs": down,
           "frac_of_link": max(up / link["h2d_gbs_alone"], down / link["d2h_gbs_alone"]),
           "link": link}
    if "d2h_gbs_mix_4_up_20_down" in link and h2d_bytes * 5 == d2h_bytes:
        res["frac_of_link_same_mix"] = down / link["d2h_gbs_mix_4_up_20_down"]   # against copies of the same mix
    return res


def e2e_dem_grid(ctx, models, dems, lines, n_job, link, steps=None, bands=None):
    """The metric end to end the way an orthorectifier calls it: a float32 DEM raster in PINNED host
    memory up (4 B/pt), line / samp / status down to pinned host arrays (20 B/pt), ground points
    generated in the kernel.  One call per image (band).  bands: this job's row bands, one per rank
    (default: equal); the result carries each rank's own seconds per step (`own_seconds`) so that the
    caller can size the bands to the rates the ranks' host links sustain."""
    import numpy as np
    import torch
    import ctypes as C
    from usgscsm_b200 import _capi
    args, barrier, max_over_ranks = ctx["args"], ctx["barrier"], ctx["max_over_ranks"]
    world, rank = ctx["world"], ctx["rank"]
    from usgscsm_b200 import sharding
    first, band = bands[rank] if bands else sharding.band_for_rank(lines, rank, world)
    nb = band * N_SAMPLES
    host_h = dems[0].samples.cpu().pin_memory()            # the same heights serve every image
    hl = torch.empty(nb, dtype=torch.float64).pin_memory()
    hs = torch.empty(nb, dtype=torch.float64).pin_memory()
    hst = torch.empty(nb, dtype=torch.int32).pin_memory()
    cdems = []
    for d in dems:
        c = _capi.Dem()
        C.memmove(C.byref(c), C.byref(d.c), C.sizeof(c))
        c.samples = host_h.data_ptr()
        cdems.append(c)

    def call():
        for m, c in zip(models, cdems):
            _capi.check(_capi.lib.usgscsm_b200_ground_to_image_dem_grid(
                m._h, C.byref(c), first, 0, band, N_SAMPLES, PREC, hl.data_ptr(), hs.data_ptr(), None,
                hst.data_ptr()))

    steps = steps or args.steps
    for _ in range(max(1, min(args.warmup, 2)) if len(models) == 1 else 1):
        call()
    barrier()
    t = time.perf_counter()
    for _ in range(steps):
        call()
    torch.cuda.synchronize()
    own = time.perf_counter() - t
    dt = max_over_ranks(own)
    barrier()
    pts_rank = nb * len(models)
    res = {"value": n_job * steps / dt / 1e6, "unit": "Mpoints/s",
           "h2d_bytes_per_step": 4 * pts_rank, "d2h_bytes_per_step": 20 * pts_rank,
           "ms_per_step": 1e3 * dt / steps, "steps": steps, "own_seconds": own / steps, "band_lines": band,
           "api": "usgscsm_b200_ground_to_image_dem_grid (host pointers, pinned; float32 DEM raster in, "
                  "line / samp / status out; ground points generated in the kernel)"}
    res.update(link_fraction(4 * pts_rank * steps, 20 * pts_rank * steps, dt, link))
    return res


def e2e_explicit(args, model, x, y, z, n, link):
    """The same metric with explicit ground-point arrays (24 B/pt up, 20 B/pt down), pinned."""
    import torch
    from usgscsm_b200 import _capi
    hx, hy, hz = (t.cpu().pin_memory() for t in (x, y, z))
    hl = torch.empty(n, dtype=torch.float64).pin_memory()
    hs = torch.empty(n, dtype=torch.float64).pin_memory()
    hst = torch.empty(n, dtype=torch.int32).pin_memory()

    def call():
        _capi.check(_capi.lib.usgscsm_b200_ground_to_image(
            model._h, n, hx.data_ptr(), hy.data_ptr(), hz.data_ptr(), PREC, hl.data_ptr(), hs.data_ptr(),
            None, hst.data_ptr()))

    call()
    steps = max(1, args.steps // 2)
    t = time.perf_counter()
    for _ in range(steps):
        call()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t
    res = {"value": n * steps / dt / 1e6, "unit": "Mpoints/s", "h2d_bytes_per_step": 24 * n,
           "d2h_bytes_per_step": 20 * n, "ms_per_step": 1e3 * dt / steps,
           "api": "usgscsm_b200_ground_to_image (host pointers, pinned)"}
    res.update(link_fraction(24 * n * steps, 20 * n * steps, dt, link))
    del res["link"]
    return res


def e2e_i2g_grid(args, model, h, lines, link):
    """Dense-grid imageToGround end to end: per-pixel heights up from pinned host memory, ground
    points and status down, pixel coordinates generated in the kernel (28 B/pt down, 8 up)."""
    import torch
    from usgscsm_b200 import _capi
    n = lines * N_SAMPLES
    hh = h[:n].cpu().pin_memory()
    hx, hy, hz = (torch.empty(n, dtype=torch.float64).pin_memory() for _ in range(3))
    hst = torch.empty(n, dtype=torch.int32).pin_memory()

    def call():
        _capi.check(_capi.lib.usgscsm_b200_image_to_ground_grid(
            model._h, 0.5, 0.5, 1.0, 1.0, lines, N_SAMPLES, hh.data_ptr(), 1, PREC,
            hx.data_ptr(), hy.data_ptr(), hz.data_ptr(), None, hst.data_ptr()))

    call()
    steps = max(1, args.steps // 2)
    t = time.perf_counter()
    for _ in range(steps):
        call()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t
    res = {"value": n * steps / dt / 1e6, "unit": "Mpoints/s", "h2d_bytes_per_step": 8 * n,
           "d2h_bytes_per_step": 28 * n, "ms_per_step": 1e3 * dt / steps,
           "api": "usgscsm_b200_image_to_ground_grid (host pointers, pinned)"}
    res.update(link_fraction(8 * n * steps, 28 * n * steps, dt, link))
    del res["link"]
    return res


def e2e_pageable(model, x, y, z, m):
    """What a numpy caller gets: ordinary (pageable) arrays in and out, staged through the
    engine's own page-locked ring by copier threads (csrc/pipeline.cuh)."""
    import numpy as np
    from usgscsm_b200 import _capi
    hx, hy, hz = (t[:m].cpu().numpy().copy() for t in (x, y, z))
    # written once before the timed call: numpy.zeros hands out untouched (calloc) pages and
    # their first-touch faults would be timed with the call
    ol, os_, ost = np.full(m, -1.0), np.full(m, -1.0), np.full(m, 7, dtype=np.uint32)

    def call(n):
        _capi.check(_capi.lib.usgscsm_b200_ground_to_image(
            model._h, n, hx.ctypes.data, hy.ctypes.data, hz.ctypes.data, PREC, ol.ctypes.data, os_.ctypes.data,
            None, ost.ctypes.data))

    call(min(m, 1 << 23))     # every staging slot allocated
    t = time.perf_counter()
    call(m)
    dt = time.perf_counter() - t
    t = time.perf_counter()
    model.groundToImageBatch(hx, hy, hz, PREC, achieved=False)
    dt_fresh = time.perf_counter() - t
    return {"value": m / dt / 1e6, "unit": "Mpoints/s", "points": m,
            "with_fresh_output_arrays": m / dt_fresh / 1e6,
            "api": "usgscsm_b200_ground_to_image (host pointers, pageable numpy arrays in and out; "
                   "with_fresh_output_arrays: outputs allocated by numpy.empty inside the timed call, "
                   "first-touch page faults included)"}


# ---- N > 1: config C5 ----------------------------------------------------------------------------
def run_c5(ctx):
    """64 line-scan images x 250 M points, images outer / row bands inner over the ranks."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import usgscsm_b200 as ub
    from usgscsm_b200 import _capi, sharding
    args, dev, local, rank, world = ctx["args"], ctx["dev"], ctx["local"], ctx["rank"], ctx["world"]
    barrier, max_over_ranks, sum_over_ranks = ctx["barrier"], ctx["max_over_ranks"], ctx["sum_over_ranks"]
    gather_floats = ctx["gather_floats"]
    lines, images = args.lines, args.images
    first, band = sharding.band_for_rank(lines, rank, world)
    nb = band * N_SAMPLES                       # points of this rank's band of one image
    n_job = images * lines * N_SAMPLES          # points of the whole job per step
    plugin = ub.UsgsAstroPlugin(device=local)
    t_build = time.perf_counter()
    models = [plugin.constructModelFromState(workload_state(1 + i, lines)) for i in range(images)]
    t_build = time.perf_counter() - t_build
    stream = torch.cuda.current_stream(dev).cuda_stream
    lib = _capi.lib
    p = lambda t: t.data_ptr()

    # heights of this band (the same synthetic DEM under every image)
    h = torch.empty(nb, dtype=torch.float64, device=dev)
    cols = torch.arange(N_SAMPLES, device=dev, dtype=torch.float64) + 0.5
    cterm = torch.cos(2 * np.pi * cols / 1024.0)
    for r0 in range(0, band, 4096):
        r1 = min(band, r0 + 4096)
        rows = torch.arange(first + r0, first + r1, device=dev, dtype=torch.float64) + 0.5
        h[r0 * N_SAMPLES:r1 * N_SAMPLES] = (500.0 * torch.sin(2 * np.pi * rows / 4096.0)[:, None] * cterm[None, :]).reshape(-1)

    # Ground points of as many images as fit (24 B/pt) are resident when their pass is timed; the
    # outputs (20 B/pt) of a pass stay sharded on the rank when they fit too, else a ring of buffers.
    free, _ = torch.cuda.mem_get_info(dev)
    budget = int(free * 0.85)
    per_pass = max(1, min(images, budget // (nb * 44)))
    if per_pass < images:
        per_pass = max(1, min(images, (budget - 4 * nb * 20) // (nb * 24)))
        ring = 4
    else:
        ring = per_pass
    passes = [list(range(s, min(images, s + per_pass))) for s in range(0, images, per_pass)]
    gx = torch.empty(per_pass * nb, dtype=torch.float64, device=dev)
    gy = torch.empty(per_pass * nb, dtype=torch.float64, device=dev)
    gz = torch.empty(per_pass * nb, dtype=torch.float64, device=dev)
    ol = torch.empty(ring * nb, dtype=torch.float64, device=dev)
    os_ = torch.empty(ring * nb, dtype=torch.float64, device=dev)
    ost = torch.empty(ring * nb, dtype=torch.int32, device=dev)
    st_g = torch.empty(nb, dtype=torch.int32, device=dev)

    def fill(pass_images):
        for k, i in enumerate(pass_images):
            _capi.check(lib.usgscsm_b200_image_to_ground_grid_dev(
                models[i]._h, first + 0.5, 0.5, 1.0, 1.0, band, N_SAMPLES, p(h), 1, PREC,
                p(gx) + 8 * k * nb, p(gy) + 8 * k * nb, p(gz) + 8 * k * nb, None, p(st_g), stream))

    def project_one(k, i):
        r = k % ring
        _capi.check(lib.usgscsm_b200_ground_to_image_dev(
            models[i]._h, nb, p(gx) + 8 * k * nb, p(gy) + 8 * k * nb, p(gz) + 8 * k * nb, PREC,
            p(ol) + 8 * r * nb, p(os_) + 8 * r * nb, None, p(ost) + 4 * r * nb, stream))

    def project(pass_images):
        for k, i in enumerate(pass_images):
            project_one(k, i)

    def step(timed):
        """One pass over the 64-image job.  Returns the timed milliseconds (max over ranks): the
        groundToImage launches only; filling the inputs of a pass (imageToGround) is outside."""
        total = 0.0
        for pi in passes:
            fill(pi)
            barrier()
            if timed:
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                project(pi)
                b.record()
                torch.cuda.synchronize()
                own = a.elapsed_time(b)
                own_ms[0] += own
                total += max_over_ranks(own)
            else:
                project(pi)
                torch.cuda.synchronize()
            barrier()
        return total

    own_ms = [0.0]
    for _ in range(max(1, args.warmup // 3)):     # a warm-up step is 64 launches per rank
        step(False)
    sampler = ClockSampler(local)   # every rank samples its own GPU; rank 0's goes to "clocks"
    sampler.start()
    launches0 = lib.usgscsm_b200_kernel_launches()
    total_ms = 0.0
    for _ in range(args.steps):
        total_ms += step(True)
    launches = lib.usgscsm_b200_kernel_launches() - launches0
    clocks = sampler.stop()
    value = n_job * args.steps / (total_ms * 1e-3) / 1e6
    # every rank's own timed milliseconds and clock, for the record
    mine = torch.tensor([own_ms[0] / args.steps, float(clocks.get("sm_mhz") or 0.0)], dtype=torch.float64, device=dev)
    everyone = [torch.empty_like(mine) for _ in range(world)]
    if world > 1:
        dist.all_gather(everyone, mine)
    else:
        everyone = [mine]
    per_rank = [{"rank": r, "ms_per_step": float(t[0]), "sm_mhz": float(t[1])} for r, t in enumerate(everyone)]

    # one more, untimed, walk over the job with an event pair and the evaluation count per image:
    # the images differ (orbit phase, attitude), and so does the quality of their projective seed
    per_image_ms, per_image_evals = [], []
    for pi in passes:
        fill(pi)
        torch.cuda.synchronize()
        for k, i in enumerate(pi):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            project_one(k, i)
            b.record()
            r = k % ring
            ev = ((ost[r * nb:(r + 1) * nb] >> _capi.ST_EVALS_SHIFT) & 0xff).double().mean()
            torch.cuda.synchronize()
            per_image_ms.append(a.elapsed_time(b))
            per_image_evals.append(float(ev.item()))
    per_launch = sorted(per_image_ms)
    mean_evals_job = sum(per_image_evals) / len(per_image_evals)

    # check: the last image's band closes the round trip
    last = (len(passes[-1]) - 1) % ring
    rows = (torch.arange(first, first + band, device=dev, dtype=torch.float64) + 0.5)[:, None]
    err = max(float((ol[last * nb:(last + 1) * nb].view(band, N_SAMPLES) - rows).abs().max()),
              float((os_[last * nb:(last + 1) * nb].view(band, N_SAMPLES) - cols[None, :]).abs().max()))
    err = max_over_ranks(err)
    flags = int(sum_over_ranks(float(((ost[last * nb:(last + 1) * nb] & _capi.ST_REFERENCE_MASK) != 0).sum().item())))
    mean_evals = sum_over_ranks(mean_evals_job) / world

    # the gather of one image's outputs (line, samp, status) onto every rank, timed on its own
    def gather_once():
        sharding.gather_bands(ol[last * nb:(last + 1) * nb], lines, N_SAMPLES)
        sharding.gather_bands(os_[last * nb:(last + 1) * nb], lines, N_SAMPLES)
        sharding.gather_bands(ost[last * nb:(last + 1) * nb], lines, N_SAMPLES)
    gather_once()
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    gather_once()
    b.record()
    torch.cuda.synchronize()
    gather_ms = max_over_ranks(a.elapsed_time(b))
    barrier()

    fp64 = measure_fp64(local) if rank == 0 else None
    result = None
    ops = opcounts()
    lin = ops["C2_linescan_nac"]["ground_to_image_model"]
    flops_pt = lin["flops_per_evaluation"] * mean_evals + lin["flops_fixed"]
    if rank == 0:
        kern_ms = total_ms / args.steps / images     # mean duration of one band launch (max over ranks)
        result = {
            "metric": METRIC, "value": value, "unit": "Mpoints/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": workload_config(args), "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": roofline_object(
                kern_ms, nb, flops_pt,
                "profiles/r2_opcount.json: %.1f per evaluation x %.3f evaluations/pt + %.1f fixed"
                % (lin["flops_per_evaluation"], mean_evals, lin["flops_fixed"]), fp64, clocks),
            "check": {"round_trip_max_px": err, "points_with_reference_flags_last_image": flags,
                      "evaluations_per_point": mean_evals},
            "job": {"images": images, "points_per_image": lines * N_SAMPLES, "band_lines_rank0": band,
                    "images_resident_per_pass": per_pass, "passes_per_step": len(passes),
                    "outputs_kept_sharded": ring == per_pass,
                    "timed_region": "the groundToImage launches of every pass, CUDA events, max over ranks per "
                                    "pass, summed; the imageToGround that produces a pass's inputs is not timed",
                    "model_build_seconds_rank0": t_build, "per_rank": per_rank,
                    "band_launch_ms_rank0": {"min": per_launch[0], "median": per_launch[len(per_launch) // 2],
                                             "max": per_launch[-1], "launches": len(per_launch),
                                             "sum": sum(per_launch)},
                    "evaluations_per_point_by_image_rank0": {"min": min(per_image_evals), "max": max(per_image_evals),
                                                             "mean": mean_evals_job}},
            "gather": {"ms_per_image": gather_ms, "bytes_per_image": 20 * lines * N_SAMPLES,
                       "gbs_per_rank_received": 20 * lines * N_SAMPLES / (gather_ms * 1e-3) / 1e9,
                       "ms_for_all_images": gather_ms * images,
                       "how": "torch.distributed.all_gather (NCCL) of line, samp and status of one image onto "
                              "every rank (sharding.gather_bands); not part of value"},
            "host": ctx["host"],
        }

    if not args.no_e2e:
        link = measure_link(dev, barrier, nbytes=1 << 29)
        agg = {k: sum_over_ranks(v) for k, v in link.items()}
        e2e_images = list(range(0, images, max(1, images // 8)))[:8]   # 8 of the 64 images, every rank
        del gx, gy, gz, ol, os_, ost
        torch.cuda.empty_cache()
        # every image gets its own footprint grid; the float32 heights are shared
        dem0 = footprint_dem(models[e2e_images[0]], lines, N_SAMPLES, dev)
        full = [dem0] + [ub.Dem(dem0.samples, footprint_grid(models[i], lines, N_SAMPLES, dev))
                         for i in e2e_images[1:]]
        e2e_models = [models[i] for i in e2e_images]
        n_e2e = len(e2e_images) * lines * N_SAMPLES
        # Row bands sized to what each rank's host link sustains: the ranks of one box share the
        # host's PCIe complex unevenly (per-rank rates under load are in link_per_rank), and with
        # equal bands the job waits for the slowest link.  Five passes with the current bands give
        # each rank's rate; the timed run uses the bands proportional to them.
        bands = sharding.row_bands(lines, world)
        history = []
        for _ in range(5 if world > 1 else 0):
            probe = e2e_dem_grid(ctx, e2e_models, full, lines, n_e2e, link, steps=1, bands=bands)
            rates = gather_floats(bands[rank][1] / probe["own_seconds"])
            history.append({"band_lines": [b for _, b in bands], "value": probe["value"]})
            bands = sharding.weighted_row_bands(lines, rates)
        e = e2e_dem_grid(ctx, e2e_models, full, lines, n_e2e, link, steps=max(1, min(2, args.steps)), bands=bands)
        e["band_lines_by_rank"] = [b for _, b in bands]
        e["seconds_per_step_by_rank"] = gather_floats(e["own_seconds"])   # how even the sized bands finish
        e["band_sizing_passes"] = history
        e["link_per_rank"] = {k: gather_floats(v) for k, v in link.items()}
        e["images_per_step"] = len(e2e_images)
        e["link_all_ranks"] = agg
        e["frac_of_box_link"] = max(sum_over_ranks(e["h2d_gbs"]) / agg["h2d_gbs_alone"],
                                    sum_over_ranks(e["d2h_gbs"]) / agg["d2h_gbs_alone"])
        # against what the box sustains when every rank only COPIES the same 4-up / 20-down mix
        e["frac_of_box_link_same_mix"] = sum_over_ranks(e["d2h_gbs"]) / agg["d2h_gbs_mix_4_up_20_down"]
        # ... and when every rank keeps copying for the same window of time (the sustained ceiling)
        e["frac_of_box_link_sustained"] = sum_over_ranks(e["d2h_gbs"]) / agg["d2h_gbs_mix_sustained"]
        if rank == 0:
            result["e2e"] = e
    return result


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
