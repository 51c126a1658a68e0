"""What the box's PCIe link gives a copy engine and an SM: pinned-memory H2D / D2H copies of the sizes
a cfg2 step moves, alone and both directions at once.  Prints GB/s."""
import torch

dev = torch.device("cuda:0")
for mb in (1.3, 10, 20.4):
    n = int(mb * 1e6) // 4
    h_in = torch.empty(n, dtype=torch.float32).pin_memory()
    h_out = torch.empty(n, dtype=torch.float32).pin_memory()
    d_a = torch.empty(n, dtype=torch.float32, device=dev)
    d_b = torch.ones(n, dtype=torch.float32, device=dev)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    def timed(fn, reps=20):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps
    t_h2d = timed(lambda: d_a.copy_(h_in, non_blocking=True))
    t_d2h = timed(lambda: h_out.copy_(d_b, non_blocking=True))
    def both():
        with torch.cuda.stream(s1):
            d_a.copy_(h_in, non_blocking=True)
        with torch.cuda.stream(s2):
            h_out.copy_(d_b, non_blocking=True)
    t_both = timed(both)
    print(f"{mb:5.1f} MB: H2D {mb / t_h2d:6.1f} GB/s ({t_h2d * 1e3:6.0f} us)   D2H {mb / t_d2h:6.1f} GB/s ({t_d2h * 1e3:6.0f} us)   "
          f"both at once {t_both * 1e3:6.0f} us per pair", flush=True)
