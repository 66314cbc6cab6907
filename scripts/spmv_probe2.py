#!/usr/bin/env python
"""Cycle accounting inside the streaming SpMV (needs a library built with B200SPAI_PROBE=1; development tool)."""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from learningsparsepreconditioner4gpu_b200 import _lib  # noqa: E402
from learningsparsepreconditioner4gpu_b200.device import DeviceCSR  # noqa: E402


def main():
    g = int(sys.argv[1]) if len(sys.argv) > 1 else 216
    n, per = g ** 3, 15
    dev = "cuda"
    lib = _lib.load()
    rows = torch.arange(n, device=dev, dtype=torch.int64)
    indptr = (torch.arange(n + 1, device=dev, dtype=torch.int64) * per).to(torch.int32)
    vals = torch.rand(n * per, device=dev, dtype=torch.float32).double()
    x = torch.randn(n, device=dev, dtype=torch.float64)
    y = torch.empty_like(x)
    offs = torch.tensor([-g * g - g - 1, -g * g - g, -g * g, -g - 1, -g, -1, 0, 1, g, g + 1, g * g, g * g + g, g * g + g + 1, g * g - 1, -g * g + 1],
                        device=dev).sort().values
    cols = (rows[:, None] + offs[None, :]).clamp_(0, n - 1).sort(dim=1).values
    A = DeviceCSR.from_torch(indptr, cols.reshape(-1).to(torch.int32), vals, (n, n))
    for _ in range(3):
        A.matvec(x, out=y, kernel=2)
    torch.cuda.synchronize()
    buf = (C.c_ulonglong * 8)()
    lib.b200spai_debug_probe.argtypes = [C.c_void_p, C.c_int]
    lib.b200spai_debug_probe(None, 1)
    reps = 10
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        A.matvec(x, out=y, kernel=2)
    e1.record()
    torch.cuda.synchronize()
    lib.b200spai_debug_probe(buf, 0)
    blocks = buf[3]
    print(json.dumps({"cfg": os.environ.get("B200SPAI_STREAM_CFG", "0"), "ms": e0.elapsed_time(e1) / reps, "blocks_per_launch": blocks / reps,
                      "consumer_wait_full_cyc_per_block": buf[0] / blocks, "consumer_process_cyc_per_block": buf[1] / blocks,
                      "producer_wait_empty_cyc_per_block": buf[2] / blocks}))


if __name__ == "__main__":
    main()
