"""Worker for tests/test_distributed_gloo.py: one rank of a world_size-N gloo group on CPU.
The band renderer is the ORACLE (this is a test: the host-side sharding logic is what is checked)."""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.dirname(HERE), HERE]

import numpy as np
import torch.distributed as dist

import helpers as H
import oracle_py as orc
from image_stitch_b200 import distributed as D


def oracle_band(plan, sources, y0, y1):
    tiles = []
    for p in D.placements_for_band(plan, y0, y1):
        s = sources[p.image_idx]
        h = s.header
        pix = (orc.decode_scanlines(np.frombuffer(bytes(s.data), np.uint8), h.width, h.height, h.bit_depth, h.color_type, s.interlace)
               if s.filtered else np.frombuffer(bytes(s.data), np.uint8))
        tiles.append(dict(width=h.width, height=h.height, bit_depth=h.bit_depth, color_type=h.color_type, pixels=pix,
                          dst_x=p.dst_x, dst_y=p.dst_y, src_off_x=p.src_off_x, src_off_y=p.src_off_y, vis_w=p.vis_w,
                          vis_h=p.vis_h, z_index=p.z_index, image_idx=p.image_idx))
    if plan.output_format == "jpeg":
        raw = orc.render_rows(plan.canvas_w, plan.canvas_h, 8, plan.background, plan.alpha_blend, tiles, y0, y1, 1)
        return orc.jpeg_encode_rows(raw.reshape(y1 - y0, plan.canvas_w * 4), plan.canvas_w, plan.canvas_h, y0, y1, plan.jpeg_quality)
    return orc.render_rows(plan.canvas_w, plan.canvas_h, plan.out_bit_depth, plan.background, plan.alpha_blend, tiles, y0, y1, 0).tobytes()


def main():
    case, out_path = sys.argv[1], sys.argv[2]
    dist.init_process_group("gloo")
    if os.environ.get("STITCH_DIST_GPU"):
        import torch
        os.environ["STITCH_B200_DEVICE"] = str(int(os.environ.get("LOCAL_RANK", "0")) % max(1, torch.cuda.device_count()))
    ins = [H.suite_bytes(n) for n in ("basn2c08", "basn6a08", "basn0g08", "basn4a08", "basi6a08", "basn2c08")]
    if case == "grid_png":
        opts = {"inputs": ins, "layout": {"columns": 2}}
    elif case == "positioned_png":
        opts = {"inputs": [{"x": 7 * i, "y": 13 * i, "source": s} for i, s in enumerate(ins)], "layout": {}}
    else:
        opts = {"inputs": ins, "layout": {"columns": 2}, "output_format": "jpeg", "jpeg_quality": 90}
    render = None if os.environ.get("STITCH_DIST_GPU") else oracle_band   # GPU test: the product's own gpu_band
    if os.environ.get("STITCH_DIST_DEVICE_DEFLATE"):
        opts["device_deflate"] = True
    if os.environ.get("STITCH_DIST_FILE"):   # the whole product path: pieces -> rank 0 -> container
        data = D.concat_to_buffer_distributed(opts, render)
        if dist.get_rank() == 0:
            with open(out_path, "wb") as f:
                f.write(data)
            with open(out_path + ".json", "w") as f:
                json.dump({}, f)
        dist.barrier()
        dist.destroy_process_group()
        return
    plan, (y0, y1), band = D.stitch_rank_band(opts, render)
    bands = D.gather_bytes(band, 0)
    if dist.get_rank() == 0:
        with open(out_path, "wb") as f:
            f.write(b"".join(bands))
        with open(out_path + ".json", "w") as f:
            json.dump({"rows": D.rank_rows(plan.canvas_h, dist.get_world_size()), "h": plan.canvas_h, "w": plan.canvas_w}, f)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
