import torch, time
n = 92405760
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(15853568, dtype=torch.uint8).pin_memory()
d2 = torch.empty(15853568, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(both):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s1):
        e0.record()
        for _ in range(20):
            d.copy_(h, non_blocking=True)
        e1.record()
    if both:
        with torch.cuda.stream(s2):
            for _ in range(20):
                h2.copy_(d2, non_blocking=True)
    torch.cuda.synchronize()
    return n * 20 / (e0.elapsed_time(e1) * 1e-3) / 1e9
for both in (False, True, False, True):
    print("h2d GB/s", round(run(both), 2), "with concurrent d2h" if both else "alone")
