"""Experiment (GPU box host): how fast can the box's cores narrow the int64 training form (x, y, polarity,
sample_index: 32 B/event) to the 5 B/event storage form in pinned memory?  Decides whether an e2e path
that narrows on the host and uploads 9 B/event can beat the plain 36 B/event upload (PCIe, ~55 GB/s)."""
import json
import os
import time

import torch

if __name__ == '__main__':
    n = int(os.environ.get('N_EVENTS', 40_000_000))
    torch.set_num_threads(len(os.sched_getaffinity(0)))
    g = torch.Generator().manual_seed(1)
    x = torch.randint(0, 256, (n,), generator=g).pin_memory()
    y = torch.randint(0, 256, (n,), generator=g).pin_memory()
    p = (torch.randint(0, 2, (n,), generator=g) * 2 - 1).pin_memory()
    b = torch.arange(8).repeat_interleave(n // 8).pin_memory()
    x16 = torch.empty(n, dtype=torch.int16).pin_memory()
    y16 = torch.empty(n, dtype=torch.int16).pin_memory()
    p8 = torch.empty(n, dtype=torch.uint8).pin_memory()
    res = {'threads': torch.get_num_threads(), 'events': n}

    def narrow():
        x16.copy_(x)
        y16.copy_(y)
        torch.gt(p, 0, out=p8.view(torch.bool))
        ok = bool((b[1:] >= b[:-1]).all())
        return ok

    for _ in range(2):
        narrow()
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        narrow()
    dt = (time.perf_counter() - t0) / reps
    res['narrow_ms'] = dt * 1e3
    res['narrow_mev_s'] = n / dt / 1e6
    res['host_GBs_read'] = 32 * n / dt / 1e9
    # plain upload of the int64 form for comparison
    dev = torch.device('cuda', 0)
    dx = torch.empty(n, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        for src in (x, y, p, b):
            dx.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    res['upload_int64_GBs'] = 3 * 4 * 8 * n / (time.perf_counter() - t0) / 1e9
    print(json.dumps(res))
