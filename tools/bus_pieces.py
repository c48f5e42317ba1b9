"""Plain pinned copies both ways at once in pieces of a given size (1 GiB per direction, one GPU): what the piece size
costs.  python tools/bus_pieces.py"""
import time

import torch

dev = torch.device("cuda", 0)
n = 1 << 30
h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True); h_in.fill_(1)
h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True); h_out.fill_(0)
d_in = torch.empty(n, dtype=torch.uint8, device=dev)
d_out = torch.empty(n, dtype=torch.uint8, device=dev); d_out.fill_(2)
s_up, s_down = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
for up_mib, down_mib in ((4, 4), (8, 8), (16, 16), (32, 32), (64, 64), (16, 4), (64, 4), (1024, 1024)):
    best = 1e9
    for _ in range(4):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        cu, cd = up_mib << 20, down_mib << 20
        ou = od = 0
        while ou < n or od < n:   # interleave the issue order by bytes
            if ou <= od and ou < n:
                with torch.cuda.stream(s_up):
                    d_in[ou:ou + cu].copy_(h_in[ou:ou + cu], non_blocking=True)
                ou += cu
            else:
                with torch.cuda.stream(s_down):
                    h_out[od:od + cd].copy_(d_out[od:od + cd], non_blocking=True)
                od += cd
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    print(f"up {up_mib:5d} MiB pieces, down {down_mib:5d} MiB pieces: {best * 1e3:6.2f} ms  ({n / best / 1e9:5.1f} GB/s per direction)", flush=True)
