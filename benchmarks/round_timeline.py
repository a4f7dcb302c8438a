#!/usr/bin/env python
"""Device-side timeline of headline rounds, from the kernels' own %globaltimer stamps.

No nsys in the image, so the picture of a round is assembled from what the kernels record anyway:
every gradient kernel stamps its message's completion (arrival record), the decoder stamps each
consumed arrival (worker_timeset) and the end of the round (timeset), all relative to the round
start on the decoder's clock (per-GPU offsets removed).

    python benchmarks/round_timeline.py [--rounds 8] [--decoder concurrent|inline] [--out FILE]
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 benchmarks/round_timeline.py

Prints one block per round: for each worker the completion time of its kernel, completion + delay,
the time the decoder consumed it (or '-' if the round ended first), and the round end; then the
split  compute (last needed completion) / wait past it / decode + push.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import torch  # noqa: E402

from erasurehead_b200 import Engine, RunConfig, triple_from_scheme  # noqa: E402
from erasurehead_b200.parallel.context import init_distributed  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rounds", type=int, default=8)
    ap.add_argument("--rows", type=int, default=524288)
    ap.add_argument("--cols", type=int, default=4096)
    ap.add_argument("--delay-scale", type=float, default=2e-3)
    ap.add_argument("--decoder", default="concurrent")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    ctx = init_distributed()
    N = ctx.world
    n, s, k = 2 * N, 1, max(1, N)
    ic, pp, cv = triple_from_scheme("approx")
    cfg = RunConfig(n + 1, n * a.rows, a.cols, "", 0, "x", ic, s, pp, cv, k, 1, "AGD", rounds=a.rounds, synthetic=True,
                    seed=1, delay_scale=a.delay_scale, quiet=True, decoder=a.decoder)
    eng = Engine(cfg, None, ctx=ctx)
    res = eng.run()
    if ctx.is_root:
        dl = eng.delays[:a.rounds]
        recs = []
        for r in range(a.rounds):
            done, arr, end = res.done[r], res.worker_timeset[r], res.timeset[r]
            used = arr >= 0
            last_needed_done = float(done[used].max()) if used.any() else float("nan")
            last_arr = float(arr[used].max()) if used.any() else float("nan")
            rec = {"round": r, "gpus": N, "decoder": a.decoder, "round_ms": end * 1e3,
                   "compute_until_last_needed_message_ms": last_needed_done * 1e3,
                   "wait_past_compute_ms": (last_arr - last_needed_done) * 1e3,
                   "decode_update_push_ms": (end - last_arr) * 1e3,
                   "workers": [{"w": w, "gpu": w % N, "done_ms": float(done[w]) * 1e3 if done[w] >= 0 else None,
                                "delay_ms": float(dl[r, w]) * 1e3,
                                "consumed_ms": float(arr[w]) * 1e3 if arr[w] >= 0 else None} for w in range(n)]}
            recs.append(rec)
            print(f"round {r}: end {end * 1e3:7.3f} ms | compute {rec['compute_until_last_needed_message_ms']:6.3f}"
                  f" + wait {rec['wait_past_compute_ms']:6.3f} + decode/push {rec['decode_update_push_ms']:6.3f}")
            for wk in rec["workers"]:
                d_ = "   -   " if wk["done_ms"] is None else f"{wk['done_ms']:7.3f}"
                rdy = "   -   " if wk["done_ms"] is None else f"{wk['done_ms'] + wk['delay_ms']:7.3f}"
                c_ = "   -   " if wk["consumed_ms"] is None else f"{wk['consumed_ms']:7.3f}"
                print(f"    worker {wk['w']:2d} (gpu {wk['gpu']}): kernel done {d_}  +delay {wk['delay_ms']:6.3f} = ready {rdy}  consumed {c_}")
        if a.out:
            json.dump(recs, open(a.out, "w"), indent=1)
    eng.close()
    if N > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
