"""Host<->device copy rates for the inference pipeline's chunk sizes (pinned memory, one GPU)."""
import time
import torch
dev = torch.device("cuda")
for rows in (16384, 65536):
    xh = torch.randn(rows, 34).pin_memory()
    oh = torch.empty(rows, 51).pin_memory()
    xd, od = torch.empty(rows, 34, device=dev), torch.randn(rows, 51, device=dev)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def run(up, down, n=50):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n):
            if up:
                with torch.cuda.stream(s1):
                    xd.copy_(xh, non_blocking=True)
            if down:
                with torch.cuda.stream(s2):
                    oh.copy_(od, non_blocking=True)
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / n * 1e6
    for _ in range(3):
        run(True, True, 5)
    u, d, b = run(True, False), run(False, True), run(True, True)
    print(f"rows {rows}: H2D {xh.numel() * 4 / 1e6:.1f} MB {u:.0f} us ({xh.numel() * 4 / u / 1e3:.1f} GB/s), "
          f"D2H {oh.numel() * 4 / 1e6:.1f} MB {d:.0f} us ({oh.numel() * 4 / d / 1e3:.1f} GB/s), both {b:.0f} us")
