"""Host->device copy bandwidth from pinned memory: one stream vs several concurrent streams (development aid)."""
import torch
dev = torch.device("cuda:0")
mb = 201
n = mb * 1024 * 1024 // 4
h = torch.empty(n, dtype=torch.float32).pin_memory()
h.fill_(1.0)
d = torch.empty(n, dtype=torch.float32, device=dev)
for k in (1, 2, 3, 4, 8):
    for chunks_per_stream in (1, 4):
        streams = [torch.cuda.Stream() for _ in range(k)]
        pieces = k * chunks_per_stream
        step = (n + pieces - 1) // pieces
        def run():
            for i in range(pieces):
                with torch.cuda.stream(streams[i % k]):
                    d[i * step:(i + 1) * step].copy_(h[i * step:(i + 1) * step], non_blocking=True)
        for _ in range(2): run()
        torch.cuda.synchronize()
        a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        import time
        t0 = time.perf_counter()
        for _ in range(10): run()
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) / 10 * 1e3
        print("%d streams x %d chunks: %.3f ms  %.1f GB/s" % (k, chunks_per_stream, ms, n * 4 / ms / 1e6))
