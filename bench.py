#!/usr/bin/env python3
"""bench.py -- BASELINE.json metric: composited output MPix/s (+ HBM GB/s vs roofline).

A "step" is one pass of the hot path over one batch of synthetic input:
BASELINE config 2 -- a 16x16 grid of 1024x1024 RGBA8 tiles (inflated, filtered
scanline streams) -> 16384x16384 filtered PNG scanlines: kernel (a) unfilter,
kernels (b)+(c) composite + filter selection.

  value : whole-job output MPix/s with the inflated streams already resident in HBM
  e2e   : same metric through the C ABI with HOST buffers (pinned): H2D of every
          tile stream and D2H of the filtered rows inside the timed region
  N > 1 : one process per GPU (torchrun); the canvas grows to 16384 x (16384*N) and
          every rank owns one row band and the tiles under it -- no collective on the
          data path ("weak" scaling)
  --impl reference : the CPU restatement of the reference (oracle/, "port": the
          reference is TypeScript and no JS runtime exists here) on all host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

TILE = 1024
GRID = 16
CANVAS = TILE * GRID            # 16384
N_DISTINCT = 8                  # distinct tile contents (re-used cyclically; every tile has its own device buffer)
ALG_BYTES_PER_PX = 8.0          # fused composite+filter: read 4*coverage(=1) + write 4
UNF_BYTES_PER_PX = 8.0          # unfilter: read 4 + write 4 per tile pixel


WORKLOAD_CONFIG = {
    "workload": "C2: 16x16 grid of 1024x1024 RGBA8 tiles (inflated filtered streams) -> 16384x16384 "
                "filtered PNG scanlines per GPU; N GPUs = N row bands of one 16384x(16384*N) canvas",
    "l2": "inputs (1.07 GB) and outputs (1.07 GB) per step exceed the 126 MB L2; no flush needed",
    "step": "unfilter(256 tiles) + composite + filter-select",
}


def profiled_traffic_bytes(summary="profiles/r1c_prof_fused.summary.csv"):
    """dram__bytes_read.sum + dram__bytes_write.sum of the composite+filter kernel from the committed
    `ncu --set full` capture of this same workload (one launch).  None when the summary is missing."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), summary)
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    got = {}
    try:
        with open(path) as f:
            for line in f:
                parts = line.strip().split(",")
                if len(parts) == 3 and parts[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum") and parts[0] not in got:
                    got[parts[0]] = float(parts[2]) * scale.get(parts[1], 1.0)
    except OSError:
        return None
    return sum(got.values()) if len(got) == 2 else None


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-sample-rows", type=int, default=1024)
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="CPU work in the cpu_baseline leg")
    ap.add_argument("--cpu-threads", type=int, default=0, help="0 = all host cores (at most one per tile column)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def make_tiles(seed: int):
    """N_DISTINCT synthetic tiles -> (raw pixels [h, w*4], inflated filtered stream)."""
    from image_stitch_b200 import synth
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(N_DISTINCT):
        pix = synth.smooth_rgba8(rng, TILE, TILE, "opaque")
        out.append((pix, synth.encode_filtered(pix, 4)))
    return out


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for k, nme in enumerate(names):
                if f[3 + k].lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---------------------------------------------------------------------------
# CPU arm: the oracle (test infrastructure) timed on the host cores.  This is the
# one place bench.py executes oracle/ -- as the baseline, never as the product.
# ---------------------------------------------------------------------------
def cpu_sample(tiles, rows: int, threads: int, repeats: int = 1):
    """Runs the reference pipeline (unfilter -> normalise -> place -> filter-select) on the CPU
    for the top `rows` canvas rows of the workload (all 16 tile columns). Returns MPix/s."""
    import oracle_py as orc
    rows = min(rows, TILE)
    streams = [tiles[c % N_DISTINCT][1] for c in range(GRID)]
    row_stride = TILE * 4 + 1

    def unfilter(c: int):
        return orc.decode_scanlines(streams[c][: rows * row_stride], TILE, rows, 8, 6, 0)

    def render(otiles, k: int, n_bands: int):
        a, b = rows * k // n_bands, rows * (k + 1) // n_bands
        if b > a:   # a band recomputes its 1-row predictor halo, like the GPU path
            orc.render_rows(CANVAS, rows, 8, bytes(4), False, otiles, a, b, 0)
        return CANVAS * (b - a)

    from concurrent.futures import ThreadPoolExecutor
    t0 = time.perf_counter()
    px = 0
    for _ in range(repeats):
        with ThreadPoolExecutor(max(1, threads)) as pool:   # ctypes releases the GIL inside the oracle
            pixels = list(pool.map(unfilter, range(GRID)))  # phase 1: one dependency chain per tile
            otiles = [dict(width=TILE, height=rows, bit_depth=8, color_type=6, pixels=pixels[c], dst_x=c * TILE,
                           dst_y=0, image_idx=c) for c in range(GRID)]
            px += sum(pool.map(lambda k: render(otiles, k, threads), range(threads)))  # phase 2: row bands
    dt = time.perf_counter() - t0
    return px / dt / 1e6, dt


def cpu_threads(args) -> int:
    return max(1, min(args.cpu_threads or (os.cpu_count() or 1), GRID))


def host_codec_rates(tiles, level: int = 6):
    """zlib inflate / deflate (level 6) on the host, outside the path's boundary (reported on
    its own line): one tile's filtered stream, one thread."""
    import zlib
    raw = tiles[0][1].tobytes()
    t0 = time.perf_counter()
    comp = zlib.compress(raw, level)
    t1 = time.perf_counter()
    back = zlib.decompress(comp)
    t2 = time.perf_counter()
    assert back == raw
    return {"deflate_MBps": len(raw) / (t1 - t0) / 1e6, "inflate_MBps": len(raw) / (t2 - t1) / 1e6, "level": level,
            "ratio": len(raw) / len(comp), "sample": f"one {TILE}x{TILE} RGBA8 tile stream ({len(raw) / 1e6:.1f} MB), 1 thread"}


def run_reference(args, rank: int, world: int):
    if rank != 0:
        return
    tiles = make_tiles(2)
    threads = cpu_threads(args)
    rows = args.cpu_sample_rows
    # One step = `reps` passes over the sample; reps is fixed before the timed region so every step is the same work.
    _, dt1 = cpu_sample(tiles, rows, threads)
    reps = max(1, int(round(2.0 / max(dt1, 1e-3))))
    for _ in range(args.warmup):
        cpu_sample(tiles, rows, threads, 1)
    t0 = time.perf_counter()
    px = 0.0
    for _ in range(args.steps):
        mp, dt = cpu_sample(tiles, rows, threads, reps)
        px += mp * dt
    total = time.perf_counter() - t0
    value = px / total
    sample = (f"{reps} x top {min(rows, TILE)} canvas rows of the 16384-wide grid per step ({GRID} tile columns), "
              f"unfilter+composite+filter, {threads} threads")
    line = {
        "impl": "reference", "metric": "composited output MPix/s", "value": value, "unit": "MPix/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": total / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": dict(WORKLOAD_CONFIG),   # identical to the GPU arm's config (the driver compares them)
        "note": "CPU restatement of the TypeScript reference (oracle/, C, -O2) on the host cores; no JS runtime on this "
                "box; each step is a bounded sample",
        "cpu_baseline": {"value": value, "unit": "MPix/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "MPix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------
def add_band_tiles(job, streams, rank: int, device_ptrs=None):
    """The 256 tiles under this rank's row band of the 16384 x (16384*N) canvas."""
    y_base = rank * CANVAS
    k = 0
    for r in range(GRID):
        for c in range(GRID):
            s = streams[k]
            kw = dict(width=TILE, height=TILE, bit_depth=8, color_type=6, interlace=0, filtered=1, dst_x=c * TILE,
                      dst_y=y_base + r * TILE, z_index=0, image_idx=rank * GRID * GRID + k)
            if device_ptrs is not None:
                job.add_tile(None, device_ptr=device_ptrs[k], length=len(s), **kw)
            else:
                job.add_tile(s, length=len(s), **kw)
            k += 1


def run_ours(args, rank: int, world: int, local_rank: int):
    import torch
    import torch.distributed as dist
    from image_stitch_b200 import _native

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- libstitch_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    tiles = make_tiles(2 + rank)
    n_tiles = GRID * GRID
    stream_len = TILE * (TILE * 4 + 1)
    canvas_h = CANVAS * world
    y0, y1 = rank * CANVAS, (rank + 1) * CANVAS
    out_row = CANVAS * 4 + 1

    # pinned host copies of every tile stream (e2e source) and of the output band (e2e sink)
    host_in = torch.empty((n_tiles, stream_len), dtype=torch.uint8, pin_memory=True)
    for k in range(n_tiles):
        host_in[k].copy_(torch.from_numpy(tiles[k % N_DISTINCT][1]))
    host_out = torch.empty(out_row * CANVAS, dtype=torch.uint8, pin_memory=True)
    host_streams = [host_in[k] for k in range(n_tiles)]

    ctx = _native.Context(local_rank)
    ts = torch.cuda.Stream(dev)      # the lib launches on this stream; torch events below are recorded on it
    ctx.set_stream(ts.cuda_stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- value: inputs resident in HBM ------------------------------------------------
    dev_in = host_in.to(dev, non_blocking=False)
    dev_ptrs = [dev_in[k].data_ptr() for k in range(n_tiles)]
    job = _native.Job(ctx, CANVAS, canvas_h, 8, _native.OUT_PNG_FILTERED, False, bytes(4), band=(y0, y1))
    add_band_tiles(job, host_streams, rank, dev_ptrs)
    launches = 0
    dec_ms, cmp_ms = [], []

    def step_resident():
        nonlocal launches
        job.invalidate()   # unfilter again: one stream submission (unfilter, plan, composite+filter), one sync
        _, t2 = job.run_band(y0, y1, download=False)
        launches += t2["launches"]
        return t2["decode_ms"], t2["compose_ms"]

    # The timed region is a few tens of milliseconds, shorter than nvidia-smi takes to start: the sampler runs
    # from before the warm-up, the warm-up continues until its first line is in, and the same load is kept up
    # after the timed region until two more lines arrive, so the samples bracket the region under load.
    sampler = ClockSampler(local_rank)
    sampler.start()
    t_w = time.perf_counter()
    n_w = 0
    while n_w < max(args.warmup, 3) or (sampler.proc and not sampler.lines and time.perf_counter() - t_w < 5.0):
        step_resident()
        n_w += 1
    launches = 0
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ts)
    for _ in range(args.steps):
        d, c = step_resident()
        dec_ms.append(d)
        cmp_ms.append(c)
    e1.record(ts)
    barrier()
    value_launches = launches
    n_lines, t_w = len(sampler.lines), time.perf_counter()
    while sampler.proc and len(sampler.lines) < n_lines + 2 and time.perf_counter() - t_w < 2.0:
        step_resident()
    torch.cuda.synchronize(dev)
    clocks = sampler.stop()
    ms_total = e0.elapsed_time(e1)
    job.close()

    # ---- e2e: host buffers in, host buffer out, every step --------------------------------
    def step_e2e():
        j = _native.Job(ctx, CANVAS, canvas_h, 8, _native.OUT_PNG_FILTERED, False, bytes(4), band=(y0, y1))
        add_band_tiles(j, host_streams, rank)
        out, tm = j.run_band(y0, y1, out=host_out)
        j.close()
        return tm

    for _ in range(2):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    e2e_tm = None
    e2e_steps = max(2, min(args.steps, 5))
    for _ in range(e2e_steps):
        e2e_tm = step_e2e()
    torch.cuda.synchronize(dev)
    e2e_s = time.perf_counter() - t0

    # ---- the same with DEFLATE on the device: compressed bytes come down the bus ---------
    with _native.Job(ctx, CANVAS, canvas_h, 8, _native.OUT_PNG_DEFLATE, False, bytes(4), band=(y0, y1)) as zj:
        zjob_bytes = zj.band_bytes(y0, y1)
    host_z = torch.empty(zjob_bytes, dtype=torch.uint8, pin_memory=True)
    z_len = 0

    def step_e2e_deflate():
        nonlocal z_len
        j = _native.Job(ctx, CANVAS, canvas_h, 8, _native.OUT_PNG_DEFLATE, False, bytes(4), band=(y0, y1))
        add_band_tiles(j, host_streams, rank)
        out, tm = j.run_band(y0, y1, out=host_z)
        z_len = len(out)
        j.close()
        return tm

    step_e2e_deflate()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e_deflate()
    torch.cuda.synchronize(dev)
    e2e_z_s = time.perf_counter() - t0

    # ---- and with INFLATE on the device as well: zlib streams up the bus, deflate blocks down ---
    import zlib
    z_distinct = [zlib.compress(bytes(tiles[k][1]), 6) for k in range(N_DISTINCT)]
    host_zin = []
    for zd in z_distinct:
        t = torch.empty(len(zd), dtype=torch.uint8, pin_memory=True)
        t.copy_(torch.frombuffer(bytearray(zd), dtype=torch.uint8))
        host_zin.append(t)
    zin_bytes = sum(len(z_distinct[k % N_DISTINCT]) for k in range(n_tiles))

    def add_zlib_tiles(j):
        k = 0
        for r in range(GRID):
            for c in range(GRID):
                zt = host_zin[k % N_DISTINCT]
                j.add_tile(zt, length=zt.numel(), width=TILE, height=TILE, bit_depth=8, color_type=6, interlace=0,
                           filtered=2, dst_x=c * TILE, dst_y=rank * CANVAS + r * TILE, z_index=0,
                           image_idx=rank * GRID * GRID + k)
                k += 1

    def step_e2e_codec():
        j = _native.Job(ctx, CANVAS, canvas_h, 8, _native.OUT_PNG_DEFLATE, False, bytes(4), band=(y0, y1))
        add_zlib_tiles(j)
        out, tm = j.run_band(y0, y1, out=host_z)
        j.close()
        return tm

    step_e2e_codec()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e_codec()
    torch.cuda.synchronize(dev)
    e2e_c_s = time.perf_counter() - t0
    # the inflate kernel alone, inputs resident
    with _native.Job(ctx, CANVAS, canvas_h, 8, _native.OUT_PNG_FILTERED, False, bytes(4), band=(y0, y1)) as ij:
        add_zlib_tiles(ij)
        ij.inflate()
        inf_ms = min(ij.inflate(force=True)["decode_ms"] for _ in range(3))

    times = torch.tensor([ms_total, e2e_s * 1e3, e2e_z_s * 1e3, e2e_c_s * 1e3, inf_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    ms_total, e2e_ms, e2e_z_ms, e2e_c_ms, inf_ms = (float(times[i]) for i in range(5))

    if rank != 0:
        ctx.close()
        return
    px_step = CANVAS * CANVAS * world / 1e6       # MPix per step, all ranks
    ms_step = ms_total / args.steps
    value = px_step / (ms_step / 1e3)
    e2e_value = px_step / (e2e_ms / e2e_steps / 1e3)
    peak, peak_src = measured_peak_gbs()
    cmp_avg, dec_avg = float(np.mean(cmp_ms)), float(np.mean(dec_ms))
    alg_bytes = ALG_BYTES_PER_PX * CANVAS * CANVAS
    achieved = alg_bytes / (cmp_avg / 1e3) / 1e9
    unf_achieved = UNF_BYTES_PER_PX * CANVAS * CANVAS / (dec_avg / 1e3) / 1e9
    line = {
        "metric": "composited output MPix/s", "value": value, "unit": "MPix/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": dict(WORKLOAD_CONFIG),
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "MPix/s", "h2d_bytes_per_step": n_tiles * stream_len * world,
                "d2h_bytes_per_step": out_row * CANVAS * world, "ms_per_step": e2e_ms / e2e_steps,
                "last_step_device_ms": e2e_tm},
        "e2e_device_deflate": {"value": px_step / (e2e_z_ms / e2e_steps / 1e3), "unit": "MPix/s",
                               "h2d_bytes_per_step": n_tiles * stream_len * world, "d2h_bytes_per_step": z_len * world,
                               "ms_per_step": e2e_z_ms / e2e_steps, "ratio": out_row * CANVAS / max(1, z_len),
                               "note": "same step, output as zlib-compatible deflate blocks made on the device "
                                       "(STITCH_OUT_PNG_DEFLATE); not the headline e2e, which stays filtered scanlines"},
        "e2e_device_codec": {"value": px_step / (e2e_c_ms / e2e_steps / 1e3), "unit": "MPix/s",
                             "h2d_bytes_per_step": zin_bytes * world, "d2h_bytes_per_step": z_len * world,
                             "ms_per_step": e2e_c_ms / e2e_steps,
                             "note": "same step with the tiles uploaded as their zlib streams (host zlib level 6) and "
                                     "inflated on the device (filtered = 2, SURVEY 8f.2), output deflated on the device: "
                                     "compressed bytes both ways; not the headline e2e"},
        "gpu_launches": value_launches,
        "roofline": {"bound": "hbm", "kernel": "composite+filter-select (kernels b+c)", "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": profiled_traffic_bytes(), "peak_source": peak_src,
                     "traffic_source": "profiles/r1c_prof_fused.summary.csv (ncu --set full, one launch, bytes)",
                     "algorithmic_bytes_per_launch": alg_bytes, "avg_ms": cmp_avg,
                     "frac_of_8TBps": achieved / 8000.0},
        "kernels": {"unfilter": {"avg_ms": dec_avg, "achieved_GBps": unf_achieved, "frac": unf_achieved / peak},
                    "compose_filter": {"avg_ms": cmp_avg, "achieved_GBps": achieved, "frac": achieved / peak},
                    "inflate": {"avg_ms": inf_ms, "inflated_GBps": n_tiles * stream_len / (inf_ms / 1e3) / 1e9,
                                "streams": n_tiles, "note": "one warp per zlib stream; serial inside a stream, "
                                "so bounded by the warp's instruction latency, not by HBM"}},
    }
    if not args.no_cpu_baseline:
        threads = cpu_threads(args)
        _, dt1 = cpu_sample(tiles, args.cpu_sample_rows, threads)
        reps = max(1, int(round(args.cpu_seconds / max(dt1, 1e-3))))
        mp, dt = cpu_sample(tiles, args.cpu_sample_rows, threads, reps)
        mp1, dt1s = cpu_sample(tiles, min(args.cpu_sample_rows, 256), 1, 1)   # the reference's own shape: one thread
        line["cpu_baseline"] = {"value": mp, "unit": "MPix/s", "cores": threads, "kind": "port",
                                "sample": f"{reps} x top {min(args.cpu_sample_rows, TILE)} rows of the same 16384-wide grid "
                                          f"({dt:.1f} s wall, {threads} threads) through the oracle "
                                          f"(C restatement of the TS reference, oracle/)",
                                "single_thread_value": mp1,
                                "single_thread_sample": f"top {min(args.cpu_sample_rows, 256)} rows, {dt1s:.1f} s"}
        line["host_codec"] = host_codec_rates(tiles)
    print(json.dumps(line), flush=True)
    ctx.close()


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world > 1:
        import torch
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        run_ours(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
