#!/usr/bin/env python
"""NVLink side of the fused gradient kernels: the same kernel writing its message into local memory
versus into the root GPU's peer-mapped slot (2 GPUs; rank 1 is the worker, rank 0 only owns the slot).

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 benchmarks/ship_bench.py
"""
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from erasurehead_b200.ops import grad, synth  # noqa: E402
from erasurehead_b200.parallel.context import init_distributed  # noqa: E402
from erasurehead_b200.parallel.workspace import Workspace  # noqa: E402


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def main():
    ctx = init_distributed()
    assert ctx.world == 2, "run with 2 GPUs"
    out = []
    for d, rows, kind in ((4096, 65536, "dense"), (8192, 32768, "dense"), (241915, 26215, "ell"), (1 << 20, 4096, "ell")):
        ws = Workspace(ctx, 2, 1, d)
        lay = ws.lay
        slot_off = int(lay.off_slots)
        ctx.barrier()
        if ctx.rank == 1:
            if kind == "dense":
                X, y = synth.synth_gmm(rows, d, 1, 0)
                beta = torch.randn(d, device="cuda") * 0.01
                sc = grad.dense_scratch(X.stride(0), "cuda")
                local = torch.empty(d, device="cuda")
                def run(ptr):
                    job = grad.make_dense_job([X], [y], beta.data_ptr(), ptr, d, [1.0], 0)
                    grad.launch_dense(job, sc)
            else:
                nnz = 44
                idx = torch.randint(0, d, (rows, nnz), device="cuda", dtype=torch.int32)
                y = torch.where(torch.rand(rows, device="cuda") > 0.5, 1.0, -1.0)
                beta = torch.randn(d, device="cuda") * 0.01
                sc = grad.ell_scratch(d, "cuda")
                local = torch.empty(d, device="cuda")
                def run(ptr):
                    job = grad.make_ell_job([idx], [y], beta.data_ptr(), ptr, d, [1.0], 0)
                    grad.launch_ell(job, sc)
            t_local = timeit(lambda: run(local.data_ptr()))
            t_peer = timeit(lambda: run(ws.root + slot_off))
            rec = {"kernel": f"grad_{kind}", "d": d, "rows": rows, "message_bytes": d * 4, "ms_local_out": t_local,
                   "ms_peer_out": t_peer, "ship_overhead_us": (t_peer - t_local) * 1e3,
                   "nvlink_ideal_us_at_770GBps": d * 4 / 770e3}
            print(json.dumps(rec), flush=True)
            out.append(rec)
        ctx.barrier()
        ws.close()
    if ctx.rank == 1 and len(sys.argv) > 1:
        json.dump(out, open(sys.argv[1], "w"), indent=1)
    torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
