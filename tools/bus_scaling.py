"""What the node's host<->device path can do with N GPUs copying at once: plain pinned cudaMemcpyAsync, no kernels.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/bus_scaling.py

Every rank owns one GPU, 1 GiB pinned in and out; all ranks start together (barrier) and the slowest rank's time
counts, like the bench's e2e.  Prints one JSON line: per direction and for both directions at once, the aggregate GB/s
(bytes per direction) and per-rank spread.  The e2e step of bench.py moves 1.07 GB up and 1.07 GB down per GPU, so
`both` is its floor: e2e_ms >= 1.0738 GB / (both_GBps_per_direction / N)."""
import json
import os
import sys
import time

import torch
import torch.distributed as dist


def main():
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    n = 1 << 30
    piece = 4 << 20   # the library uploads tile by tile (4 MiB) and downloads in pieces of that size
    h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True); h_in.fill_(1)
    h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True); h_out.fill_(0)
    d_in = torch.empty(n, dtype=torch.uint8, device=dev)
    d_out = torch.empty(n, dtype=torch.uint8, device=dev); d_out.fill_(2)
    s_up, s_down = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()

    def run(up, down, pieces):
        c = piece if pieces else n
        barrier()
        t0 = time.perf_counter()
        for o in range(0, n, c):
            if up:
                with torch.cuda.stream(s_up):
                    d_in[o:o + c].copy_(h_in[o:o + c], non_blocking=True)
            if down:
                with torch.cuda.stream(s_down):
                    h_out[o:o + c].copy_(d_out[o:o + c], non_blocking=True)
        torch.cuda.synchronize(dev)
        dt = time.perf_counter() - t0
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            tmin = t.clone(); dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
            return float(tmax), float(tmin)
        return dt, dt

    res = {}
    for name, up, down, pieces in (("h2d", 1, 0, False), ("d2h", 0, 1, False), ("both", 1, 1, False), ("both_4MiB_pieces", 1, 1, True)):
        best = None
        for _ in range(4):
            tmax, tmin = run(up, down, pieces)
            if best is None or tmax < best[0]:
                best = (tmax, tmin)
        res[name] = {"ms_slowest_rank": best[0] * 1e3, "ms_fastest_rank": best[1] * 1e3,
                     "aggregate_GBps_per_direction": world * n / best[0] / 1e9, "per_gpu_GBps_per_direction": n / best[0] / 1e9}
    if rank == 0:
        aff = sorted(os.sched_getaffinity(0))
        print(json.dumps({"n_gpus": world, "bytes_per_direction_per_gpu": n, "host_cpus_allowed": len(aff),
                          "numa_nodes_visible": len([d for d in os.listdir("/sys/devices/system/node") if d.startswith("node")])
                          if os.path.isdir("/sys/devices/system/node") else None, **res}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
