"""Pinned host -> device bandwidth of the box for the transfer shapes of bench.py's end-to-end leg."""
import time, torch
n = 8192; npx = 640 * 480
hg = torch.empty((n, npx), dtype=torch.uint8).pin_memory(); hd = torch.empty((n, npx), dtype=torch.int16).pin_memory()
dg = torch.empty_like(hg, device="cuda"); dd = torch.empty_like(hd, device="cuda")
s = torch.cuda.Stream()
def run(label, fn, nbytes):
    for _ in range(2): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(3): fn()
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 3
    print(f"{label}: {nbytes / dt / 1e9:.1f} GB/s ({dt * 1e3:.1f} ms)")
with torch.cuda.stream(s):
    run("gray 2.5 GB one copy", lambda: dg.copy_(hg, non_blocking=True), hg.numel())
    run("depth 5.0 GB one copy", lambda: dd.copy_(hd, non_blocking=True), hd.numel() * 2)
    run("both", lambda: (dg.copy_(hg, non_blocking=True), dd.copy_(hd, non_blocking=True)), hg.numel() * 3)
    q = n // 8
    run("both, 8 chunks each", lambda: [(dg[i*q:(i+1)*q].copy_(hg[i*q:(i+1)*q], non_blocking=True), dd[i*q:(i+1)*q].copy_(hd[i*q:(i+1)*q], non_blocking=True)) for i in range(8)], hg.numel() * 3)
