#!/usr/bin/env python
"""Host->device copy rate from pinned memory: one large copy vs a stream of batch-sized copies
(314 KB fp32 / 78 KB uint8 MNIST batches).  Explains the end-to-end bound of the host-fed loop."""
import json
import time

import torch

dev = torch.device("cuda:0")
out = {}
for name, nbytes, reps in (("64MB", 64 << 20, 20), ("313600B", 313600, 2000), ("78400B", 78400, 2000)):
    src = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(4)]
    dst = [torch.empty(nbytes, dtype=torch.uint8, device=dev) for _ in range(4)]
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for i in range(8):
            dst[i % 4].copy_(src[i % 4], non_blocking=True)
        st.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(st)
        for i in range(reps):
            dst[i % 4].copy_(src[i % 4], non_blocking=True)
        e1.record(st)
        t_issue = time.perf_counter() - t0
        st.synchronize()
    ms = e0.elapsed_time(e1)
    out[name] = {"GB_per_s": nbytes * reps / ms / 1e6, "us_per_copy": ms * 1e3 / reps,
                 "host_issue_us_per_copy": t_issue * 1e6 / reps}
print(json.dumps(out))
