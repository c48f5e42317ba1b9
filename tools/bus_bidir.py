import torch, time
dev = torch.device("cuda", 0)
N = 1 << 30
h_in = torch.empty(N, dtype=torch.uint8, pin_memory=True); h_in.fill_(1)
h_out = torch.empty(N, dtype=torch.uint8, pin_memory=True)
d_in = torch.empty(N, dtype=torch.uint8, device=dev)
d_out = torch.empty(N, dtype=torch.uint8, device=dev); d_out.fill_(2)
s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
def run(up, down, chunks=1):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    c = N // chunks
    for k in range(chunks):
        if up:
            with torch.cuda.stream(s1): d_in[k*c:(k+1)*c].copy_(h_in[k*c:(k+1)*c], non_blocking=True)
        if down:
            with torch.cuda.stream(s2): h_out[k*c:(k+1)*c].copy_(d_out[k*c:(k+1)*c], non_blocking=True)
    torch.cuda.synchronize()
    return time.perf_counter() - t0
for name, u, d, ch in (("H2D", 1, 0, 1), ("D2H", 0, 1, 1), ("both", 1, 1, 1), ("both x16 chunks", 1, 1, 16), ("H2D x256 chunks", 1, 0, 256)):
    ts = [run(u, d, ch) for _ in range(4)]
    t = min(ts)
    print(f"{name}: {t*1e3:.1f} ms  {(u+d)*N/t/1e9:.1f} GB/s aggregate")

# ---- the library's upload pattern: 256 copies of 4 MiB from separate pinned buffers, an event after each ----
def run2(events, down_chunks=16):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    c = N // 256
    evs = []
    for k in range(256):
        with torch.cuda.stream(s1):
            d_in[k*c:(k+1)*c].copy_(h_in[k*c:(k+1)*c], non_blocking=True)
            if events:
                e = torch.cuda.Event(); e.record(s1); evs.append(e)
    cd = N // down_chunks
    for k in range(down_chunks):
        with torch.cuda.stream(s2):
            h_out[k*cd:(k+1)*cd].copy_(d_out[k*cd:(k+1)*cd], non_blocking=True)
    torch.cuda.synchronize()
    return time.perf_counter() - t0
for ev in (False, True):
    t = min(run2(ev) for _ in range(4))
    print(f"256 x 4 MiB up (events={ev}) + 16 x 64 MiB down: {t*1e3:.1f} ms")

ups = [torch.cuda.Stream(dev) for _ in range(4)]
def run3(n_up_streams, copy_mib=4, down_chunks=16):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = (N >> 20) // copy_mib
    c = N // n
    for k in range(n):
        with torch.cuda.stream(ups[k % n_up_streams]):
            d_in[k*c:(k+1)*c].copy_(h_in[k*c:(k+1)*c], non_blocking=True)
    cd = N // down_chunks
    for k in range(down_chunks):
        with torch.cuda.stream(s2):
            h_out[k*cd:(k+1)*cd].copy_(d_out[k*cd:(k+1)*cd], non_blocking=True)
    torch.cuda.synchronize()
    return time.perf_counter() - t0
for ns in (1, 2, 4):
    for mib in (4, 16, 64):
        t = min(run3(ns, mib) for _ in range(3))
        print(f"{ns} upload stream(s), {mib} MiB copies + 16 x 64 MiB down: {t*1e3:.1f} ms")

for mib_up, mib_down in ((4, 4), (4, 16), (4, 1)):
    def run4():
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        n = (N >> 20) // mib_up; c = N // n
        nd = (N >> 20) // mib_down; cd = N // nd
        for k in range(max(n, nd)):
            if k < n:
                with torch.cuda.stream(s1): d_in[k*c:(k+1)*c].copy_(h_in[k*c:(k+1)*c], non_blocking=True)
            for kk in range(k * nd // max(n, nd), (k + 1) * nd // max(n, nd)) if nd >= n else ([k] if k < nd else []):
                with torch.cuda.stream(s2): h_out[kk*cd:(kk+1)*cd].copy_(d_out[kk*cd:(kk+1)*cd], non_blocking=True)
        torch.cuda.synchronize()
        return time.perf_counter() - t0
    t = min(run4() for _ in range(3))
    print(f"{mib_up} MiB up copies + {mib_down} MiB down copies: {t*1e3:.1f} ms")
