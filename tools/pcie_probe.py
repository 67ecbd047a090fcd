"""PCIe probe: pinned H2D / D2H alone and concurrently (what bounds b2a_read_arrow_host).  One JSON line."""
import json
import torch

GB = 1 << 30
n = 2 * GB
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_a = torch.empty(n, dtype=torch.uint8, device="cuda")
d_b = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=3):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        s1.synchronize()
        s2.synchronize()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def h2d():
    with torch.cuda.stream(s1):
        d_a.copy_(h_in, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        h_out.copy_(d_b, non_blocking=True)


def both():
    h2d()
    d2h()


def chunks(k):
    def f():
        step = n // k
        for i in range(k):
            with torch.cuda.stream(s1):
                d_a[i * step:(i + 1) * step].copy_(h_in[i * step:(i + 1) * step], non_blocking=True)
            with torch.cuda.stream(s2):
                h_out[i * step:(i + 1) * step].copy_(d_b[i * step:(i + 1) * step], non_blocking=True)
    return f


import time
out = {}
for name, fn, nbytes in (("h2d", h2d, n), ("d2h", d2h, n), ("both", both, 2 * n), ("both_64chunks", chunks(64), 2 * n)):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    fn()
    torch.cuda.synchronize()
    w = (time.perf_counter() - t0) * 1e3
    ms = min(timed(fn), w)
    out[name] = {"ms": round(ms, 2), "GBps_total": round(nbytes / ms / 1e6, 1)}
print(json.dumps(out))

# ---- H2D in small pieces while large D2H copies run (the shape of b2a_read_arrow_host's traffic)
def contention(h2d_chunk, d2h_chunk, h2d_total=GB, d2h_total=GB + GB // 4):
    torch.cuda.synchronize()
    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a0.record(s1)
    b0.record(s2)
    i = j = 0
    # interleave the enqueues so neither stream's queue runs dry
    while i < h2d_total or j < d2h_total:
        if i < h2d_total:
            with torch.cuda.stream(s1):
                d_a[i:i + h2d_chunk].copy_(h_in[i:i + h2d_chunk], non_blocking=True)
            i += h2d_chunk
        if j < d2h_total and j * h2d_total <= i * d2h_total:
            with torch.cuda.stream(s2):
                h_out[j:j + d2h_chunk].copy_(d_b[j:j + d2h_chunk], non_blocking=True)
            j += d2h_chunk
    a1.record(s1)
    b1.record(s2)
    torch.cuda.synchronize()
    return round(h2d_total / a0.elapsed_time(a1) / 1e6, 1), round(d2h_total / b0.elapsed_time(b1) / 1e6, 1)


MB = 1 << 20
res = {}
for hc in (1, 4, 8, 32, 128):
    contention(hc * MB, 256 * MB)
    res[f"h2d_{hc}MB_vs_d2h_256MB"] = contention(hc * MB, 256 * MB)
res["h2d_4MB_vs_d2h_16MB"] = contention(4 * MB, 16 * MB)
res["h2d_4MB_alone"] = contention(4 * MB, 256 * MB, d2h_total=0)
res["h2d_1MB_alone"] = contention(1 * MB, 256 * MB, d2h_total=0)
print(json.dumps(res))

# ---- two H2D streams: does a second copy in flight hide the per-copy start-up under D2H contention?
s1b = torch.cuda.Stream()


def contention2(h2d_chunk, d2h_chunk, h2d_total=GB, d2h_total=GB + GB // 4):
    torch.cuda.synchronize()
    a0, a1, a2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a0.record(s1)
    b0.record(s2)
    i = j = k = 0
    while i < h2d_total or j < d2h_total:
        if i < h2d_total:
            with torch.cuda.stream(s1 if k % 2 == 0 else s1b):
                d_a[i:i + h2d_chunk].copy_(h_in[i:i + h2d_chunk], non_blocking=True)
            i += h2d_chunk
            k += 1
        if j < d2h_total and j * h2d_total <= i * d2h_total:
            with torch.cuda.stream(s2):
                h_out[j:j + d2h_chunk].copy_(d_b[j:j + d2h_chunk], non_blocking=True)
            j += d2h_chunk
    a1.record(s1)
    a2.record(s1b)
    b1.record(s2)
    torch.cuda.synchronize()
    t = max(a0.elapsed_time(a1), a0.elapsed_time(a2))
    return round(h2d_total / t / 1e6, 1), round(d2h_total / b0.elapsed_time(b1) / 1e6, 1)


res = {}
for hc in (1, 4, 8):
    contention2(hc * MB, 256 * MB)
    res[f"2streams_h2d_{hc}MB_vs_d2h_256MB"] = contention2(hc * MB, 256 * MB)
print(json.dumps(res))
