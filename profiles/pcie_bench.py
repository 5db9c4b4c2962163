"""Raw PCIe bandwidth of the box with pinned buffers of the e2e leg's sizes (CUDA events)."""
import torch
h2d = torch.empty(26215104 // 4, dtype=torch.float32).pin_memory()
d2h = torch.empty(16394240 // 4, dtype=torch.float32).pin_memory()
a = torch.empty_like(h2d, device="cuda"); b = torch.empty_like(d2h, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(up, down, n=50):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        if up:
            with torch.cuda.stream(s1): a.copy_(h2d, non_blocking=True)
        if down:
            with torch.cuda.stream(s2): d2h.copy_(b, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
s1.wait_stream(torch.cuda.current_stream()); s2.wait_stream(torch.cuda.current_stream())
for name, up, down in (("H2D 26.2 MB alone", 1, 0), ("D2H 16.4 MB alone", 0, 1), ("both directions", 1, 1)):
    run(up, down, 5)
    ms = run(up, down)
    gb = (26215104 * up + 16394240 * down) / ms / 1e6
    print("%-20s %.3f ms  %.1f GB/s  -> %.2f M audio-s/s if this were the whole step" % (name, ms, gb, 512 / ms / 1e3))
