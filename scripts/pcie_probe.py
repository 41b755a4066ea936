import torch, time
n = 2709504000
h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
d = torch.empty(n, dtype=torch.uint8, device='cuda')
ho = torch.empty(1007616000, dtype=torch.uint8, pin_memory=True)
do = torch.empty(1007616000, dtype=torch.uint8, device='cuda')
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for both in (0, 1):
    for rep in range(3):
        torch.cuda.synchronize(); t = time.perf_counter()
        with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
        if both:
            with torch.cuda.stream(s2): ho.copy_(do, non_blocking=True)
        torch.cuda.synchronize(); dt = time.perf_counter() - t
    print('duplex' if both else 'h2d only', 'ms %.2f' % (dt * 1e3), 'H2D GB/s %.1f' % (n / dt / 1e9))
