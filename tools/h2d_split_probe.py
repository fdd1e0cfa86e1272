import torch, time
n = 122880000
h = torch.empty(n, dtype=torch.uint8).pin_memory(); h.fill_(1)
d = torch.empty(n, dtype=torch.uint8, device="cuda")
def run(k):
    streams = [torch.cuda.Stream() for _ in range(k)]
    parts = [(i * n // k, (i + 1) * n // k) for i in range(k)]
    for _ in range(3):
        for s, (a, b) in zip(streams, parts):
            with torch.cuda.stream(s):
                d[a:b].copy_(h[a:b], non_blocking=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(30):
        for s, (a, b) in zip(streams, parts):
            with torch.cuda.stream(s):
                d[a:b].copy_(h[a:b], non_blocking=True)
    torch.cuda.synchronize()
    return 30 * n / (time.perf_counter() - t0) / 1e9
for k in (1, 2, 4, 1, 2):
    print(k, "streams: %.1f GB/s" % run(k))
