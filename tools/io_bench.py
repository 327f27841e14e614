#!/usr/bin/env python
"""Host <-> HBM bandwidth of the state I/O entry points: setAmps and the read-back, from pinned and
from pageable host memory, next to what one plain copy of the same pageable array achieves (torch .cuda()).
Writes one JSON object to stdout."""
import ctypes
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from quest_b200 import capi  # noqa: E402


def main():
    nq = int(sys.argv[1]) if len(sys.argv) > 1 else 28
    n = 1 << nq
    lib = capi.load_product()
    env = lib.createQuESTEnv()
    q = lib.createQureg(nq, env)
    dp = ctypes.POINTER(ctypes.c_double)
    page_re, page_im = np.random.rand(n), np.random.rand(n)
    pin_re = torch.from_numpy(page_re).clone().pin_memory()
    pin_im = torch.from_numpy(page_im).clone().pin_memory()
    out = {"amplitudes": n, "bytes_per_direction": 16 * n, "io_threads": os.environ.get("QUEST_B200_IO_THREADS", "default")}

    def timed(fn, reps=3):
        best = 1e9
        for _ in range(reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            best = min(best, time.perf_counter() - t0)
        return 16 * n / best / 1e9

    out["setAmps_pinned_GBps"] = timed(lambda: lib.setAmps(q, 0, ctypes.cast(pin_re.data_ptr(), dp), ctypes.cast(pin_im.data_ptr(), dp), n))
    out["setAmps_pageable_GBps"] = timed(lambda: lib.setAmps(q, 0, page_re.ctypes.data_as(dp), page_im.ctypes.data_as(dp), n))
    back_re, back_im = np.empty(n), np.empty(n)
    out["readback_pageable_GBps"] = timed(lambda: lib.qb200_copyAmpsToHost(q, 0, n, back_re.ctypes.data_as(dp), back_im.ctypes.data_as(dp)))
    assert np.array_equal(back_re, page_re) and np.array_equal(back_im, page_im)
    pb_re, pb_im = torch.empty(n, dtype=torch.float64).pin_memory(), torch.empty(n, dtype=torch.float64).pin_memory()
    out["readback_pinned_GBps"] = timed(lambda: lib.qb200_copyAmpsToHost(q, 0, n, ctypes.cast(pb_re.data_ptr(), dp), ctypes.cast(pb_im.data_ptr(), dp)))
    t_re, t_im = torch.from_numpy(page_re), torch.from_numpy(page_im)

    def plain():
        a = t_re.cuda()
        b = t_im.cuda()
        del a, b
    out["plain_copy_pageable_GBps"] = timed(plain)
    lib.destroyQureg(q, env)
    lib.destroyQuESTEnv(env)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
