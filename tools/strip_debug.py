"""Small strip-kernel scenarios, one per process (a CUDA fault is sticky): which box kind misbehaves.
  python tools/strip_debug.py            runs every scenario in a subprocess
  python tools/strip_debug.py NAME       runs one"""
import os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))

SCEN = {
    # name: (canvas_w, canvas_h, blend, [(w, h, alpha, x, y, z)])
    "bg_only": (3000, 20, False, []),
    "direct_plus_bg_right": (3000, 20, False, [(1024, 20, "opaque", 0, 0, 0)]),
    "direct_plus_bg_below": (2048, 30, False, [(2048, 10, "opaque", 0, 0, 0)]),
    "comp_narrow": (600, 20, False, [(100, 20, "opaque", 0, 0, 0)]),
    "comp_edge": (3000, 20, False, [(1000, 20, "opaque", 100, 0, 0)]),
    "unaligned_pitch": (3000, 20, False, [(1001, 20, "opaque", 0, 0, 0)]),
    "two_overlap": (3000, 40, False, [(1200, 30, "opaque", 0, 0, 0), (1200, 30, "opaque", 1000, 5, 1)]),
    "odd_width": (5003, 30, False, [(1200, 30, "opaque", 0, 0, 0), (1001, 30, "opaque", 3000, 0, 0)]),
    "blend": (3000, 40, True, [(1200, 30, "opaque", 0, 0, 0), (1200, 30, "uniform", 1000, 5, 1)]),
}

def _t2(cw, n, alphas=("opaque", "uniform", "binary")):
    rng = np.random.default_rng(22)
    spec, x = [], 3
    for i in range(6):
        w = int(rng.integers(500, 1100)); h = int(rng.integers(40, 90)); y = int(rng.integers(0, 20))
        if i < n:
            spec.append((w, h, alphas[i % len(alphas)], x, y, i))
        x += w + int(rng.integers(-40, 60))
    return spec

SCEN.update({
    "cw24": (3000, 20, False, [(1200, 20, "opaque", 1000, 0, 0)]),
    "cw22": (3000, 20, False, [(1200, 20, "opaque", 1002, 0, 0)]),
    "cw23": (3000, 20, False, [(1200, 20, "opaque", 1001, 0, 0)]),
    "cw20": (3000, 20, False, [(1200, 20, "opaque", 1004, 0, 0)]),
    "py_0_90": None, "py_17_60": None, "py_0_8": None, "py_40_90": None, "py_0_40": None,
    "t2": (5003, 90, False, _t2(5003, 6)),
    "t2_w5004": (5004, 90, False, _t2(5004, 6)),
    "t2_opaque": (5003, 90, False, _t2(5003, 6, ("opaque",))),
    "t2_first2": (5003, 90, False, _t2(5003, 2)),
    "t2_first1": (5003, 90, False, _t2(5003, 1)),
    "t2_h20": (5003, 20, False, _t2(5003, 6)),
    "t2_band": (5003, 90, False, _t2(5003, 6), [(17, 60)]),
    "t2_band_a": (5003, 90, False, _t2(5003, 6), [(17, 18)]),
    "t2_band_b": (5003, 90, False, _t2(5003, 6), [(16, 60)]),
    "t2_band_1": (5003, 90, False, _t2(5003, 1), [(17, 60)]),
    "t2_band_op": (5003, 90, False, _t2(5003, 6, ("opaque",)), [(17, 60)]),
    "t2_band_w": (5004, 90, False, _t2(5004, 6), [(17, 60)]),
})


def run_pytest_layout(blend, bands):
    """The exact layout of tests/test_gpu_strip.py::test_tiles_at_odd_offsets_and_gaps."""
    import oracle_py as orc
    from image_stitch_b200 import _native
    import test_gpu_strip as T
    rng = np.random.default_rng(22)
    cw, ch = 5003, 90
    tiles, x = [], 3
    for i in range(6):
        w = int(rng.integers(500, 1100))
        tiles.append(T._tile(rng, w, int(rng.integers(40, 90)), ["opaque", "uniform", "binary"][i % 3], x, int(rng.integers(0, 20)), i, i, cw, ch))
        x += w + int(rng.integers(-40, 60))
    for t_ in tiles:
        print("tile", t_["image_idx"], "w", t_["width"], "h", t_["height"], "x", t_["dst_x"], "y", t_["dst_y"], "vis", t_["vis_w"], t_["vis_h"], flush=True)
    ctx = _native.Context(0)
    bg = bytes([10, 200, 30, 128])
    for (a, b) in bands:
        got = T._render(ctx, cw, ch, 8, blend, bg, tiles, [(a, b)])[0]
        want = orc.render_rows(cw, ch, 8, bg, blend, [t_ for t_ in tiles if t_["vis_w"] and t_["vis_h"]], a, b, 0)
        bad = np.nonzero(got != want)[0]
        print("pytest layout", (a, b), "OK" if bad.size == 0 else f"{bad.size} differ", flush=True)


def run(name):
    if name.startswith("py_"):
        a, b = name[3:].split("_")
        return run_pytest_layout(False, [(int(a), int(b))])
    import oracle_py as orc
    from image_stitch_b200 import _native, synth
    cw, ch, blend, spec = SCEN[name][:4]
    bands = SCEN[name][4] if len(SCEN[name]) > 4 else [(0, ch)]
    rng = np.random.default_rng(5)
    tiles = []
    for i, (w, h, alpha, x, y, z) in enumerate(spec):
        tiles.append(dict(width=w, height=h, bit_depth=8, color_type=6, pixels=synth.smooth_rgba8(rng, w, h, alpha), dst_x=x, dst_y=y,
                          src_off_x=0, src_off_y=0, vis_w=min(w, cw - x), vis_h=min(h, ch - y), z_index=z, image_idx=i))
    ctx = _native.Context(0)
    bg = bytes([9, 8, 7, 200])
    with _native.Job(ctx, cw, ch, 8, _native.OUT_PNG_FILTERED, blend, bg) as job:
        for t in tiles:
            job.add_tile(t["pixels"].reshape(-1), width=t["width"], height=t["height"], bit_depth=8, color_type=6, filtered=0,
                         dst_x=t["dst_x"], dst_y=t["dst_y"], vis_w=t["vis_w"], vis_h=t["vis_h"], z_index=t["z_index"], image_idx=t["image_idx"])
        for (a, b) in bands:
            got, _ = job.run_band(a, b)
            got = np.array(got, copy=True)
            want = orc.render_rows(cw, ch, 8, bg, blend, tiles, a, b, 0)
            bad = np.nonzero(got != want)[0]
            row = cw * 4 + 1
            print(name, (a, b), "OK" if bad.size == 0 else f"{bad.size} bytes differ, first row {a + bad[0] // row} byte {bad[0] % row}", flush=True)

if __name__ == "__main__":
    if len(sys.argv) == 2:
        run(sys.argv[1])
    else:
        for n in (sys.argv[2:] if len(sys.argv) > 2 else SCEN):
            r = subprocess.run([sys.executable, __file__, n], capture_output=True, text=True, timeout=120)
            print(" ; ".join(r.stdout.strip().splitlines() or [n + " no output"]), "|", (r.stderr.strip().splitlines() or [""])[-1][:160], flush=True)
