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
C3_CANVAS, C3_TILE, C3_SIDE = 32768, 2048, 16
ALG_BYTES_PER_PX = 8.0          # fused composite+filter: read 4*coverage(=1) + write 4
UNF_BYTES_PER_PX = 8.0          # unfilter: read 4 + write 4 per tile pixel


WORKLOAD_CONFIG = {
    "workload": "C2: 16x16 grid of 1024x1024 RGBA8 tiles (inflated filtered streams, a random PNG filter type per "
                "tile row) -> 16384x16384 filtered PNG scanlines per GPU; N GPUs = N row bands of ONE 16384x(16384*N) "
                "canvas (band k also uploads and unfilters the 16 tiles above it: its first row's predictor is their last row)",
    "l2": "inputs (1.07 GB) and outputs (1.07 GB) per step exceed the 126 MB L2; no flush needed",
    "step": "unfilter(256 tiles, + 16 halo tiles on ranks > 0) + composite + filter-select",
}


def profiled_traffic_bytes(*summaries, scale: float = 1.0):
    """dram__bytes_read.sum + dram__bytes_write.sum of the profiled kernel(s) from committed `ncu --set full`
    captures of this same workload (one launch each), times `scale`.  None when a summary is missing."""
    summaries = summaries or ("profiles/r2e_prof_fused_c2.summary.csv",)
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    total = 0.0
    for summary in summaries:
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), summary)
        got = {}
        try:
            with open(path) as f:
                for line in f:
                    parts = line.strip().split(",")
                    if len(parts) == 3 and parts[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):   # every kernel of the capture
                        got[parts[0]] = got.get(parts[0], 0.0) + float(parts[2]) * unit.get(parts[1], 1.0)
        except OSError:
            return None
        if len(got) != 2:
            return None
        total += sum(got.values())
    return total * scale


def e2e_floor(world: int, h2d_per_gpu: float, d2h_per_gpu: float, e2e_ms: float):
    """What plain pinned copies achieve on this class of node with `world` GPUs copying both ways at once
    (tools/bus_scaling.py, committed as profiles/r2_bus_scaling.jsonl: the node's host side delivers 50-65 GB/s
    per direction in total, however many GPUs share it), and how close the e2e step comes to it."""
    path = os.path.join(ROOT, "profiles", "r2_bus_scaling.jsonl")
    try:
        rows = [json.loads(l) for l in open(path) if l.strip()]
    except OSError:
        return None
    row = next((r for r in rows if r["n_gpus"] == world), None)
    if row is None:
        return None
    rate = row["both"]["per_gpu_GBps_per_direction"] * 1e9
    floor_ms = max(h2d_per_gpu, d2h_per_gpu) / rate * 1e3
    return {"floor_ms_per_step": floor_ms, "e2e_over_floor": e2e_ms / floor_ms, "node_GBps_per_direction_both_ways": row["both"]["aggregate_GBps_per_direction"],
            "source": "profiles/r2_bus_scaling.jsonl (plain cudaMemcpyAsync, 1 GiB per direction per GPU, all GPUs at once, slowest rank)"}


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


def make_tiles(seed: int, force: int = 5):
    """N_DISTINCT synthetic tiles -> (raw pixels [h, w*4], inflated filtered stream).
    force: the PNG filter type of the tile rows (synth.encode_filtered): 5 = a random type per row (the headline:
    every unfilter path in one stream), 4 = Paeth everywhere, -1 = the min-sum heuristic (Sub + Average on these tiles)."""
    from image_stitch_b200 import synth
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(N_DISTINCT):
        pix = synth.smooth_rgba8(rng, TILE, TILE, "opaque")
        out.append((pix, synth.encode_filtered(pix, 4, force, rng)))
    return out


class _DevBuf:
    """A device pointer as a __cuda_array_interface__ object (for torch.as_tensor)."""
    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def band_hash(job, y0: int, dev):
    """Two position-weighted sums over the filtered rows of the band the job rendered last (device side, rows
    without their padding): equal bands <=> equal hashes for any single-byte error; used to compare a band rendered
    by its own rank with the same rows of the whole canvas rendered by rank 0."""
    import torch
    ptr, pitch, lead, rb = job.band_device()
    rows = job.last_rows
    flat = torch.as_tensor(_DevBuf(ptr, pitch * rows), device=dev)
    h1 = h2 = 0
    for a in range(0, rows, 2048):
        x = flat[a * pitch:(min(rows, a + 2048)) * pitch].view(-1, pitch)[:, lead:lead + rb]
        rs = x.sum(dim=1, dtype=torch.int64)
        cs = x.sum(dim=0, dtype=torch.int64)
        h1 += int((rs * (torch.arange(rs.numel(), device=dev, dtype=torch.int64) + (y0 + a + 1))).sum().item()) & ((1 << 62) - 1)
        h2 += int((cs * (torch.arange(cs.numel(), device=dev, dtype=torch.int64) + 1)).sum().item()) & ((1 << 62) - 1)
    return [h1 & ((1 << 62) - 1), h2 & ((1 << 62) - 1)]


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
def add_band_tiles(job, streams, rank: int, device_ptrs=None, filtered: int = 1):
    """The tiles this rank's row band of the 16384 x (16384*N) canvas needs: the 256 under it and, for rank > 0,
    the 16 of the grid row above it (streams[256:272]) -- the band's first row is filtered against their last row,
    which the rank recomputes instead of receiving.  streams: host buffers (or just lengths with
    device_ptrs)."""
    y_base = rank * CANVAS
    places = [(r, c) for r in range(GRID) for c in range(GRID)]
    if rank > 0:
        places += [(-1, c) for c in range(GRID)]
    for k, (r, c) in enumerate(places):
        s = streams[k]
        idx = rank * GRID * GRID + k if r >= 0 else (rank - 1) * GRID * GRID + (GRID - 1) * GRID + c
        kw = dict(width=TILE, height=TILE, bit_depth=8, color_type=6, interlace=0, filtered=filtered, dst_x=c * TILE,
                  dst_y=y_base + r * TILE, z_index=0, image_idx=idx)
        if device_ptrs is not None:
            job.add_tile(None, device_ptr=device_ptrs[k], length=len(s), **kw)
        else:
            job.add_tile(s, length=len(s), **kw)


def tile_index(rank: int, k: int) -> int:
    """Which of the rank's N_DISTINCT tile contents grid position k of band `rank` shows."""
    return k % N_DISTINCT


# ---------------------------------------------------------------------------
# Sections of the GPU arm beside the headline workload
# ---------------------------------------------------------------------------
def _c3_setup(dev, need):
    """Device buffers (one per tile: nothing is shared, so nothing is served from L2 by accident) for the C3 tiles
    in `need`, filled from four distinct synthetic tiles."""
    import torch
    from image_stitch_b200 import synth
    rng = np.random.default_rng(33)
    raws = [torch.from_numpy(synth.smooth_rgba8(rng, C3_TILE, C3_TILE, k).reshape(-1)).to(dev) for k in synth.C3_CONTENT]
    return {t["image_idx"]: raws[t["content"]].clone() for t in need}


def _c3_job(ctx, layout, bufs, blend: bool, band):
    from image_stitch_b200 import _native
    job = _native.Job(ctx, C3_CANVAS, C3_CANVAS, 8, _native.OUT_PNG_FILTERED, blend, bytes([255, 255, 255, 128]), band=band)
    for t in layout:
        if t["image_idx"] in bufs and t["vis_w"] and t["vis_h"]:
            b = bufs[t["image_idx"]]
            job.add_tile(None, device_ptr=b.data_ptr(), length=b.numel(), width=C3_TILE, height=C3_TILE, bit_depth=8, color_type=6,
                         filtered=0, dst_x=t["dst_x"], dst_y=t["dst_y"], src_off_x=t["src_off_x"], src_off_y=t["src_off_y"],
                         vis_w=t["vis_w"], vis_h=t["vis_h"], z_index=t["z_index"], image_idx=t["image_idx"])
    return job


def c3_section(ctx, dev, rank: int, world: int, barrier):
    """Composite + filter-select of this rank's band of the C3 canvas, blending on and off (device-timed, inputs
    resident): ms per band, and the band's hash for the seam check."""
    from image_stitch_b200 import synth
    layout = synth.c3_layout(np.random.default_rng(3), C3_CANVAS, C3_SIDE, C3_TILE)
    y0, y1 = rank * C3_CANVAS // world, (rank + 1) * C3_CANVAS // world
    need = [t for t in layout if t["vis_h"] and t["dst_y"] < y1 and t["dst_y"] + t["vis_h"] > max(0, y0 - 1)]
    bufs = _c3_setup(dev, need)
    out = {"tiles": len(need)}
    for blend in (True, False):
        job = _c3_job(ctx, layout, bufs, blend, (y0, y1))
        for _ in range(2):
            job.run_band(y0, y1, download=False)
        barrier()
        ms = [job.run_band(y0, y1, download=False)[1]["compose_ms"] for _ in range(3)]
        barrier()
        out["blend_ms" if blend else "noblend_ms"] = float(np.mean(ms))
        if blend:
            out["hash"] = band_hash(job, y0, dev)
        job.close()
    del bufs
    return out


def c3_full_check(ctx, dev, world: int, hashes) -> bool:
    """Rank 0: the whole C3 canvas as one job, band by band, against the hashes of the ranks' own bands."""
    from image_stitch_b200 import synth
    layout = synth.c3_layout(np.random.default_rng(3), C3_CANVAS, C3_SIDE, C3_TILE)
    bufs = _c3_setup(dev, [t for t in layout if t["vis_h"]])
    job = _c3_job(ctx, layout, bufs, True, (0, C3_CANVAS))
    ok = True
    for k in range(world):
        a, b = k * C3_CANVAS // world, (k + 1) * C3_CANVAS // world
        job.run_band(a, b, download=False)
        ok = ok and band_hash(job, a, dev) == list(hashes[k])
    job.close()
    return bool(ok)


def seam_check_c2(ctx, dev, world: int, hashes) -> bool:
    """Rank 0: the whole 16384 x (16384 * world) mosaic as one job (every rank's 256 tiles, no halo tiles), band by
    band, against the hashes of the bands the ranks rendered alone from their tiles plus the 16 halo tiles."""
    from image_stitch_b200 import _native
    job = _native.Job(ctx, CANVAS, CANVAS * world, 8, _native.OUT_PNG_FILTERED, False, bytes(4))
    keep = []
    for k in range(world):
        tk = make_tiles(2 + k)
        keep.append(tk)
        for i in range(GRID * GRID):
            st = tk[i % N_DISTINCT][1]
            job.add_tile(st, length=len(st), width=TILE, height=TILE, bit_depth=8, color_type=6, interlace=0, filtered=1,
                         dst_x=(i % GRID) * TILE, dst_y=k * CANVAS + (i // GRID) * TILE, z_index=0, image_idx=k * GRID * GRID + i)
    ok = True
    for k in range(world):
        job.run_band(k * CANVAS, (k + 1) * CANVAS, download=False)
        ok = ok and band_hash(job, k * CANVAS, dev) == list(hashes[k])
    job.close()
    return bool(ok)


def unfilter_variants(ctx, dev, peak: float):
    """Kernel (a) alone on the 256 tiles of C2 by filter type of the tile rows (SURVEY.md 8d asks for per-filter
    figures): Paeth everywhere, and what the reference's own heuristic picks for these tiles (Sub + Average)."""
    import torch
    from image_stitch_b200 import _native
    out = {}
    for name, force in (("paeth", 4), ("heuristic_sub_average", -1)):
        tv = make_tiles(2, force)
        dev_in = torch.stack([torch.from_numpy(tv[k % N_DISTINCT][1]) for k in range(GRID * GRID)]).to(dev)
        with _native.Job(ctx, CANVAS, CANVAS, 8, _native.OUT_PNG_FILTERED, False, bytes(4)) as job:
            add_band_tiles(job, [dev_in[k] for k in range(GRID * GRID)], 0, [dev_in[k].data_ptr() for k in range(GRID * GRID)])
            job.decode(force=True)
            ms = min(job.decode(force=True)["decode_ms"] for _ in range(4))
        gbs = UNF_BYTES_PER_PX * CANVAS * CANVAS / (ms / 1e3) / 1e9
        out[name] = {"avg_ms": ms, "achieved_GBps": gbs, "frac": gbs / peak}
        del dev_in
    return out


def jpeg_section(ctx, dev, peak: float):
    """Kernel (d): a 16x16 mosaic of 2048^2 opaque tiles -> 32768^2 baseline JPEG, q = 90, restart interval of one MCU
    row (BASELINE C5's 65536^2 exceeds SOF0's 16-bit size fields, so it cannot be a JPEG)."""
    import torch
    from image_stitch_b200 import _native, synth
    rng = np.random.default_rng(5)
    raws = [torch.from_numpy(synth.smooth_rgba8(rng, C3_TILE, C3_TILE, "opaque").reshape(-1)).to(dev) for _ in range(4)]
    bufs = [raws[i % 4].clone() for i in range(C3_SIDE * C3_SIDE)]
    res = {"workload": "16x16 mosaic of 2048x2048 RGBA8 tiles -> 32768x32768 baseline JPEG (4:4:4, q=90, DRI = one MCU row)"}
    with _native.Job(ctx, C3_CANVAS, C3_CANVAS, 8, _native.OUT_JPEG_ENTROPY, False, bytes(4), jpeg_quality=90) as job:
        for i, b in enumerate(bufs):
            job.add_tile(None, device_ptr=b.data_ptr(), length=b.numel(), width=C3_TILE, height=C3_TILE, bit_depth=8, color_type=6,
                         filtered=0, dst_x=(i % C3_SIDE) * C3_TILE, dst_y=(i // C3_SIDE) * C3_TILE, image_idx=i)
        job.run_band(0, C3_CANVAS, download=False)
        ms = [job.run_band(0, C3_CANVAS, download=False)[1]["compose_ms"] for _ in range(3)]
        os.environ["STITCH_JPEG_DIRECT"] = "0"   # the same band composited into a canvas first (what a band that is not an exact mosaic costs)
        try:
            job.run_band(0, C3_CANVAS, download=False)
            ms_canvas = [job.run_band(0, C3_CANVAS, download=False)[1]["compose_ms"] for _ in range(3)]
        finally:
            del os.environ["STITCH_JPEG_DIRECT"]
        job.run_band(0, C3_CANVAS, download=False)
        _, _, _, n_bytes = job.band_device()
        host = torch.empty(n_bytes + 64, dtype=torch.uint8, pin_memory=True)
        t0 = time.perf_counter()
        out, _ = job.run_band(0, C3_CANVAS, out=host)
        d2h_ms = (time.perf_counter() - t0) * 1e3
    px = C3_CANVAS * C3_CANVAS
    avg = float(np.mean(ms))
    alg = 4.0 * px + n_bytes
    res.update({"avg_ms": avg, "value_MPix_s": px / 1e6 / (avg / 1e3), "entropy_bytes": int(n_bytes), "bits_per_pixel": 8.0 * n_bytes / px,
                "roofline": {"bound": "hbm", "algorithmic_bytes": alg, "achieved_GBps": alg / (avg / 1e3) / 1e9,
                             "frac": alg / (avg / 1e3) / 1e9 / peak,
                             "note": "one-pass route of csrc/jpeg.cu: the MCU kernel reads the tiles in place (an exact mosaic: no canvas), colour + FDCT + "
                                     "quantise + Huffman coding in registers and shared memory, then row info + stuffing; algorithmic = 4 B/px read + "
                                     "entropy bytes written; the MCU kernel is bound by instruction issue (profiles/r2h_prof_jpeg.summary.csv), not by HBM"},
                "through_canvas": {"avg_ms": float(np.mean(ms_canvas)), "value_MPix_s": px / 1e6 / (float(np.mean(ms_canvas)) / 1e3)},
                "with_d2h_of_the_entropy_bytes_ms": d2h_ms})
    return res


def jpeg_decode_section(ctx):
    """16 baseline JPEG tiles of 2048^2 (4:2:0, q = 85, no restart markers: the common kind) decoded on the
    device, composited and filtered -- wall time from the first add_tile_jpeg (host buffers, the files cross the bus
    compressed) to the finished band on the device.  The files are made with Pillow when the box has it."""
    import io
    from image_stitch_b200 import _native, synth
    try:
        from PIL import Image
    except ImportError:
        return {"unavailable": "Pillow is not installed: no JPEG file to decode"}
    tile, grid = C3_TILE, 4
    pic = synth.smooth_rgba8(np.random.default_rng(8), tile, tile, "opaque").reshape(tile, tile, 4)[..., :3]
    buf = io.BytesIO()
    Image.fromarray(np.ascontiguousarray(pic), "RGB").save(buf, "JPEG", quality=85, subsampling=2)
    data = buf.getvalue()
    t0 = time.perf_counter()
    Image.open(io.BytesIO(data)).load()
    cpu_ms = (time.perf_counter() - t0) * 1e3
    best = None
    for _ in range(4):
        with _native.Job(ctx, tile * grid, tile * grid, 8, _native.OUT_PNG_FILTERED, False, bytes(4)) as job:
            t0 = time.perf_counter()
            for k in range(grid * grid):
                job.add_tile_jpeg(data, dst_x=(k % grid) * tile, dst_y=(k // grid) * tile, image_idx=k)
            t1 = time.perf_counter()
            _, tm = job.run_band(0, tile * grid, download=False)
            t2 = time.perf_counter()
        ms = ((t1 - t0) * 1e3, (t2 - t0) * 1e3, tm["compose_ms"])
        best = ms if best is None or ms[1] < best[1] else best
    px = grid * grid * tile * tile
    # the same tiles to ONE JPEG (q = 90): decoded on the device, read in place by the encoder, entropy bytes downloaded
    jj = None
    host = None
    for _ in range(4):
        with _native.Job(ctx, tile * grid, tile * grid, 8, _native.OUT_JPEG_ENTROPY, False, bytes(4), jpeg_quality=90) as job:
            t0 = time.perf_counter()
            for k in range(grid * grid):
                job.add_tile_jpeg(data, dst_x=(k % grid) * tile, dst_y=(k // grid) * tile, image_idx=k)
            if host is None:
                import torch
                host = torch.empty(job.band_bytes(0, 8) * (tile * grid // 8) // 8 + (1 << 20), dtype=torch.uint8, pin_memory=True)
            out, tm = job.run_band(0, tile * grid, out=host)
            ms = (time.perf_counter() - t0) * 1e3
        if jj is None or ms < jj[0]:
            jj = (ms, int(len(out)), tm["compose_ms"])
    return {"workload": f"{grid * grid} JPEG tiles of {tile}x{tile} (4:2:0, q=85, {len(data)} bytes each, no restart markers) -> "
                        f"{tile * grid}x{tile * grid} filtered PNG scanlines",
            "queued_ms": best[0], "wall_ms": best[1], "composite_filter_ms": best[2], "value_MPix_s": px / 1e6 / (best[1] / 1e3),
            "h2d_bytes": len(data) * grid * grid, "libjpeg_turbo_one_core_ms_per_tile": cpu_ms,
            "jpeg_to_jpeg": {"wall_ms": jj[0], "value_MPix_s": px / 1e6 / (jj[0] / 1e3), "d2h_bytes": jj[1], "encode_ms": jj[2],
                             "note": "the same 16 files decoded on the device and coded again as one JPEG (q = 90, the decoded tiles read "
                                     "in place by the MCU kernel), host buffers both ways"},
            "note": "wall clock around add_tile_jpeg x 16 + run_band (decode kernels are queued at add time on four streams); "
                    "byte-exact against libjpeg-turbo (tests/test_gpu_jpeg_decode.py)"}


def run_ours(args, rank: int, world: int, local_rank: int):
    import torch
    import torch.distributed as dist
    from image_stitch_b200 import _native

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- libstitch_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    tiles = make_tiles(2 + rank)
    tiles_above = make_tiles(2 + rank - 1) if rank > 0 else None
    n_own = GRID * GRID
    n_tiles = n_own + (GRID if rank > 0 else 0)      # + the halo tiles (grid row above the band)
    stream_len = TILE * (TILE * 4 + 1)
    canvas_h = CANVAS * world
    y0, y1 = rank * CANVAS, (rank + 1) * CANVAS
    out_row = CANVAS * 4 + 1

    def content(k: int):
        """(tile set, index) shown at position k of this rank's tile list (add_band_tiles order)."""
        return (tiles, k % N_DISTINCT) if k < n_own else (tiles_above, ((GRID - 1) * GRID + (k - n_own)) % N_DISTINCT)

    # pinned host copies of every tile stream (e2e source) and of the output band (e2e sink)
    host_in = torch.empty((n_tiles, stream_len), dtype=torch.uint8, pin_memory=True)
    for k in range(n_tiles):
        ts_, i_ = content(k)
        host_in[k].copy_(torch.from_numpy(ts_[i_][1]))
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
    my_hash = band_hash(job, y0, dev)     # this rank's band, rendered from its own tiles + halo tiles
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
    z_sets = {id(tiles): [zlib.compress(bytes(tiles[k][1]), 6) for k in range(N_DISTINCT)]}
    if tiles_above is not None:
        z_sets[id(tiles_above)] = [zlib.compress(bytes(tiles_above[k][1]), 6) for k in range(N_DISTINCT)]
    pinned_z = {}
    for key, zs in z_sets.items():
        for i_, zd in enumerate(zs):
            t = torch.empty(len(zd), dtype=torch.uint8, pin_memory=True)
            t.copy_(torch.frombuffer(bytearray(zd), dtype=torch.uint8))
            pinned_z[(key, i_)] = t
    host_zin = [pinned_z[(id(content(k)[0]), content(k)[1])] for k in range(n_tiles)]
    zin_bytes = sum(t.numel() for t in host_zin)

    def step_e2e_codec():
        j = _native.Job(ctx, CANVAS, canvas_h, 8, _native.OUT_PNG_DEFLATE, False, bytes(4), band=(y0, y1))
        add_band_tiles(j, host_zin, rank, filtered=2)
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
        add_band_tiles(ij, host_zin, rank, filtered=2)
        ij.inflate()
        inf_ms = min(ij.inflate(force=True)["decode_ms"] for _ in range(3))
    del dev_in
    # ---- C3 (north_star's target shape), strong-scaled: ONE 32768^2 canvas of 256 positioned 2048^2 tiles, zIndex
    #      alpha blending, cut into `world` row bands; a rank uploads the tiles that touch its band or the row above it
    c3 = c3_section(ctx, dev, rank, world, barrier)

    times = torch.tensor([ms_total, e2e_s * 1e3, e2e_z_s * 1e3, e2e_c_s * 1e3, inf_ms, c3["blend_ms"], c3["noblend_ms"]],
                         dtype=torch.float64, device=dev)
    hashes = [None] * world
    c3_hashes = [None] * world
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
        dist.all_gather_object(hashes, my_hash)
        dist.all_gather_object(c3_hashes, c3["hash"])
    ms_total, e2e_ms, e2e_z_ms, e2e_c_ms, inf_ms, c3_blend_ms, c3_noblend_ms = (float(times[i]) for i in range(7))

    if rank != 0:
        ctx.close()
        return
    seam = None
    if world > 1:   # the whole canvas as ONE job on this GPU: every band must hash like the rank that rendered it alone
        seam = {"c2_bands_equal": seam_check_c2(ctx, dev, world, hashes), "c3_bands_equal": c3_full_check(ctx, dev, world, c3_hashes),
                "method": "rank 0 renders the whole canvas as one job (no halo tiles, no bands) and compares position-weighted "
                          "row/column sums of each band's filtered bytes with those of the rank that rendered the band alone"}
    px_step = CANVAS * CANVAS * world / 1e6       # MPix per step, all ranks
    ms_step = ms_total / args.steps
    value = px_step / (ms_step / 1e3)
    e2e_value = px_step / (e2e_ms / e2e_steps / 1e3)
    peak, peak_src = measured_peak_gbs()
    cmp_avg, dec_avg = float(np.mean(cmp_ms)), float(np.mean(dec_ms))
    alg_bytes = ALG_BYTES_PER_PX * CANVAS * CANVAS
    achieved = alg_bytes / (cmp_avg / 1e3) / 1e9
    unf_achieved = UNF_BYTES_PER_PX * n_tiles * TILE * TILE / (dec_avg / 1e3) / 1e9
    h2d_all = (world * n_own + (world - 1) * GRID) * stream_len
    c3_px = C3_CANVAS * C3_CANVAS
    c3_alg = 8.0 * c3_px       # 4 B/px x coverage 1.0 read + 4 B/px written
    line = {
        "metric": "composited output MPix/s", "value": value, "unit": "MPix/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": dict(WORKLOAD_CONFIG),
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "MPix/s", "h2d_bytes_per_step": h2d_all,
                "d2h_bytes_per_step": out_row * CANVAS * world, "ms_per_step": e2e_ms / e2e_steps,
                "last_step_device_ms": e2e_tm},
        "e2e_device_deflate": {"value": px_step / (e2e_z_ms / e2e_steps / 1e3), "unit": "MPix/s",
                               "h2d_bytes_per_step": h2d_all, "d2h_bytes_per_step": z_len * world,
                               "ms_per_step": e2e_z_ms / e2e_steps, "ratio": out_row * CANVAS / max(1, z_len),
                               "note": "same step, output as zlib-compatible deflate blocks made on the device "
                                       "(STITCH_OUT_PNG_DEFLATE); not the headline e2e, which stays filtered scanlines"},
        "e2e_device_codec": {"value": px_step / (e2e_c_ms / e2e_steps / 1e3), "unit": "MPix/s",
                             "h2d_bytes_per_step": zin_bytes * world, "d2h_bytes_per_step": z_len * world,
                             "ms_per_step": e2e_c_ms / e2e_steps,
                             "note": "same step with the tiles uploaded as their zlib streams (host zlib level 6) and "
                                     "inflated on the device (filtered = 2, SURVEY 8f.2), output deflated on the device: "
                                     "compressed bytes both ways; not the headline e2e"},
        "e2e_floor": e2e_floor(world, h2d_all / world, out_row * CANVAS, e2e_ms / e2e_steps),
        "gpu_launches": value_launches,
        "roofline": {"bound": "hbm", "kernel": "composite+filter-select (kernels b+c), C2", "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": profiled_traffic_bytes(), "peak_source": peak_src,
                     "traffic_source": "profiles/r2e_prof_fused_c2.summary.csv (ncu --set full of this kernel on this workload, one launch, bytes)",
                     "algorithmic_bytes_per_launch": alg_bytes, "avg_ms": cmp_avg,
                     "frac_of_8TBps": achieved / 8000.0},
        "roofline_c3": {"bound": "hbm", "kernel": "composite + filter-select on C3; with blending: composite to a canvas in HBM, then the unit kernel "
                                                  "reads it back; without: the unit kernel's plan mode (plan + prepare + fused, no canvas)",
                        "workload": "C3: 256 positioned overlapping 2048x2048 RGBA8 tiles (half opaque, a quarter alpha 1..254, a quarter 0/255 "
                                    "masks), zIndex alpha blending, 32768x32768 canvas, background #FFFFFF80; every tile has its own buffer",
                        "avg_ms": c3_blend_ms, "achieved": c3_alg / (c3_blend_ms / 1e3) / 1e9, "peak": peak, "unit": "GB/s",
                        "frac": c3_alg / (c3_blend_ms / 1e3) / 1e9 / peak, "value_MPix_s": c3_px / 1e6 / (c3_blend_ms / 1e3),
                        "algorithmic_bytes_per_launch": c3_alg, "traffic": profiled_traffic_bytes("profiles/r2e_prof_c3_blend.summary.csv", scale=8.0),
                        "traffic_source": "8 x (compose_plan_kernel + fused_kernel of one 4096-row band), profiles/r2e_prof_c3_blend.summary.csv",
                        "blending_off": {"avg_ms": c3_noblend_ms, "frac": c3_alg / (c3_noblend_ms / 1e3) / 1e9 / peak,
                                         "value_MPix_s": c3_px / 1e6 / (c3_noblend_ms / 1e3),
                                         "traffic": profiled_traffic_bytes("profiles/r2e_prof_c3_noblend.summary.csv", scale=8.0),
                                         "traffic_source": "8 x (fused_plan_iv + fused_prep + fused_kernel of one 4096-row band), "
                                                           "profiles/r2e_prof_c3_noblend.summary.csv"},
                        "scaling": "strong: the one canvas is cut into n_gpus row bands, max over ranks", "rows_per_gpu": C3_CANVAS // world},
        "kernels": {"unfilter": {"avg_ms": dec_avg, "achieved_GBps": unf_achieved, "frac": unf_achieved / peak, "tiles": n_tiles,
                                 "filter_types": "random per row"},
                    "compose_filter": {"avg_ms": cmp_avg, "achieved_GBps": achieved, "frac": achieved / peak},
                    "inflate": {"avg_ms": inf_ms, "inflated_GBps": n_tiles * stream_len / (inf_ms / 1e3) / 1e9,
                                "streams": n_tiles, "note": "a warp pair per zlib stream; serial inside a stream, "
                                "so bounded by the warps' instruction latency, not by HBM"}},
    }
    if seam is not None:
        line["seam_check"] = seam
    if world == 1:
        line["kernels"]["unfilter_variants"] = unfilter_variants(ctx, dev, peak)
        line["jpeg"] = jpeg_section(ctx, dev, peak)
        line["jpeg_decode"] = jpeg_decode_section(ctx)
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
