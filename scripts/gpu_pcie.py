"""PCIe ceiling for the host-buffer ray path: pinned H2D of 537 MB alone, and with a concurrent 268 MB D2H."""
import torch
n_in, n_out = 16777216 * 32, 16777216 * 16
h_in = torch.empty(n_in, dtype=torch.uint8).pin_memory(); h_out = torch.empty(n_out, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n_in, dtype=torch.uint8, device="cuda"); d_out = torch.empty(n_out, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(both, reps=10):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s1.wait_event(e0); s2.wait_event(e0)
        with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
        if both:
            with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
        ev1 = torch.cuda.Event(); ev2 = torch.cuda.Event()
        ev1.record(s1); ev2.record(s2)
        torch.cuda.current_stream().wait_event(ev1); torch.cuda.current_stream().wait_event(ev2)
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
a = run(False); b = run(True)
print("H2D alone: %.2f ms = %.1f GB/s;  H2D + concurrent D2H: %.2f ms (H2D %.1f GB/s) -> ceiling %.0f Mrays/s" % (a, n_in / a / 1e6, b, n_in / b / 1e6, 16777216 / b / 1e3))
