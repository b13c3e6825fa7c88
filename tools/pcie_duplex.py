"""H2D alone, D2H alone and both at once (two streams, pinned memory): is PCIe full duplex here?"""
import time, torch
n = 1 << 30  # bytes
h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True)
d_in = torch.empty(n, dtype=torch.uint8, device='cuda')
d_out = torch.empty(n, dtype=torch.uint8, device='cuda')
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(up, down, reps=3):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        if up:
            with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
        if down:
            with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize(); best = min(best, time.perf_counter() - t0)
    return best
for name, a, b in (('h2d', 1, 0), ('d2h', 0, 1), ('both', 1, 1)):
    t = run(a, b)
    print(f'{name}: {t*1e3:.1f} ms, {(a+b)*n/t/1e9:.1f} GB/s total')
# chunked interleave on both streams (16 x 64 MB)
c = n // 16
torch.cuda.synchronize(); t0 = time.perf_counter()
for k in range(16):
    with torch.cuda.stream(s1): d_in[k*c:(k+1)*c].copy_(h_in[k*c:(k+1)*c], non_blocking=True)
    with torch.cuda.stream(s2): h_out[k*c:(k+1)*c].copy_(d_out[k*c:(k+1)*c], non_blocking=True)
torch.cuda.synchronize(); t = time.perf_counter() - t0
print(f'both chunked: {t*1e3:.1f} ms, {2*n/t/1e9:.1f} GB/s total')
