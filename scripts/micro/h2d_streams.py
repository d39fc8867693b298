"""Pinned host -> device copy rate of one 11.5 MB step input (bench.py's e2e arena) as 1, 2, 4, 8 concurrent chunks on as many
streams: tells whether a slow box is limited by the link or by one DMA queue.  python scripts/micro/h2d_streams.py"""
import time
import torch

dev = torch.device("cuda", 0)
total = 11_468_800
h = torch.empty(total, dtype=torch.uint8).pin_memory()
d = [torch.empty(total, dtype=torch.uint8, device=dev) for _ in range(2)]
for k in (1, 2, 4, 8, 1):
    streams = [torch.cuda.Stream(dev) for _ in range(k)]
    cuts = [total * i // k // 256 * 256 for i in range(k)] + [total]

    def step(i):
        for j, s in enumerate(streams):
            with torch.cuda.stream(s):
                d[i & 1][cuts[j]:cuts[j + 1]].copy_(h[cuts[j]:cuts[j + 1]], non_blocking=True)

    for i in range(5):
        step(i)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(5):
        t0 = time.perf_counter()
        for i in range(20):
            step(i)
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    print("chunks/streams %d: %.1f GB/s (%.3f ms per 11.5 MB step)" % (k, 20 * total / best / 1e9, best / 20 * 1e3), flush=True)
# device -> host for reference
hd = torch.empty(total, dtype=torch.uint8).pin_memory()
torch.cuda.synchronize(); t0 = time.perf_counter()
for i in range(20):
    hd.copy_(d[0], non_blocking=True)
torch.cuda.synchronize()
print("d2h 1 stream: %.1f GB/s" % (20 * total / (time.perf_counter() - t0) / 1e9))
