"""H2D (2.4 GB) and D2H (1.6 GB) at once on two streams, cut into N parts each, the D2H part k waiting
for the H2D part k (like the upload / K1 / download pipeline of transitions._pipelined_states):
how does the wall time depend on the number of parts?"""
import time, torch
up_bytes, down_bytes = 2_419_200_000, 1_612_800_000
h_in = torch.empty(up_bytes, dtype=torch.uint8, pin_memory=True)
h_out = torch.empty(down_bytes, dtype=torch.uint8, pin_memory=True)
d_in = torch.empty(up_bytes, dtype=torch.uint8, device='cuda')
d_out = torch.empty(down_bytes, dtype=torch.uint8, device='cuda')
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(parts, dependent):
    best = 1e9
    for _ in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for k in range(parts):
            a, b = up_bytes * k // parts, up_bytes * (k + 1) // parts
            c, d = down_bytes * k // parts, down_bytes * (k + 1) // parts
            with torch.cuda.stream(s1):
                d_in[a:b].copy_(h_in[a:b], non_blocking=True)
                ev = torch.cuda.Event(); ev.record(s1)
            with torch.cuda.stream(s2):
                if dependent:
                    s2.wait_event(ev)
                h_out[c:d].copy_(d_out[c:d], non_blocking=True)
        torch.cuda.synchronize(); best = min(best, time.perf_counter() - t0)
    return best


for parts in (1, 2, 4, 8, 16, 32, 64):
    for dep in (False, True):
        t = run(parts, dep)
        print(f'parts {parts:3d} dependent {int(dep)}: {t*1e3:6.1f} ms   H2D {up_bytes/t/1e9:5.1f} GB/s', flush=True)
