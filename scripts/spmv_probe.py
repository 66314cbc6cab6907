#!/usr/bin/env python
"""Where does the streaming SpMV lose time?  Synthetic 15-entries-per-row matrices with controlled gather locality
(development tool).  a: every column == row (gathers always hit L1);  b: a dense band row-7..row+7;  c: 3-D 7/15-point
offsets (mesh-like);  also torch's copy bandwidth on the same box."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from learningsparsepreconditioner4gpu_b200.device import DeviceCSR  # noqa: E402


def timed(fn, reps=20):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    g = int(sys.argv[1]) if len(sys.argv) > 1 else 216
    n, per = g ** 3, 15
    dev = "cuda"
    rows = torch.arange(n, device=dev, dtype=torch.int64)
    indptr = (torch.arange(n + 1, device=dev, dtype=torch.int64) * per).to(torch.int32)
    vals = torch.rand(n * per, device=dev, dtype=torch.float32).double()
    x = torch.randn(n, device=dev, dtype=torch.float64)
    y = torch.empty_like(x)
    out = {"n": n, "per_row": per, "cfg": os.environ.get("B200SPAI_STREAM_CFG", "0"), "col16": os.environ.get("B200SPAI_NO_COL16", "0") != "1"}
    a = torch.empty(1 << 28, device=dev, dtype=torch.float32)
    b = torch.empty_like(a)
    ms = timed(lambda: b.copy_(a))
    out["copy_GBs"] = round(2 * a.numel() * 4 / ms / 1e6, 1)
    del a, b
    offs_mesh = torch.tensor([-g * g - g - 1, -g * g - g, -g * g, -g - 1, -g, -1, 0, 1, g, g + 1, g * g, g * g + g, g * g + g + 1, g * g - 1, -g * g + 1],
                             device=dev).sort().values
    pats = {
        "same": torch.zeros(per, device=dev, dtype=torch.int64),
        "band": torch.arange(per, device=dev) - 7,
        "mesh": offs_mesh,
    }
    for name, offs in pats.items():
        cols = (rows[:, None] + offs[None, :]).clamp_(0, n - 1)
        if name != "same":
            cols = cols.sort(dim=1).values
        A = DeviceCSR.from_torch(indptr, cols.reshape(-1).to(torch.int32), vals, (n, n))
        ms = timed(lambda: A.matvec(x, out=y, kernel=2))
        real = n * per * (6 if out["col16"] else 8) + n * (8 if out["col16"] else 4) + 16 * n
        out[name] = {"ms": round(ms, 4), "real_GBs": round(real / ms / 1e6, 1)}
        del A, cols
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
