"""Multi-process engine check (helper, launched by torchrun from test_multi_process.py).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tests/mp_engine_check.py <device> <scheme> [<scheme> ...]

Every rank builds the same seeded dataset, runs the engine and the root compares the iterates,
arrival sets and decode weights with the float64 oracle.  Exit code 0 = all schemes agree.
"""
import os
import sys

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))

import numpy as np  # noqa: E402

from erasurehead_b200 import Engine, RunConfig, build_scheme, oracle, triple_from_scheme  # noqa: E402
from erasurehead_b200.data import generate_gmm, generate_onehot  # noqa: E402
from erasurehead_b200.parallel.context import init_distributed  # noqa: E402

KW = {
    "naive": dict(s=0), "coded": dict(s=2, seed=3), "replication": dict(s=1), "avoidstragg": dict(s=2),
    "approx": dict(s=1, k=5), "partial_coded": dict(n=4, s=1, P=4, rows=12 * 160, seed=2),
    "partial_replication": dict(n=4, s=1, P=4, rows=12 * 160), "sparse": dict(s=1, k=6, rows=8 * 120, d=700),
    "softmax": dict(s=1, k=6, rows=8 * 256, d=96, model="softmax", classes=5, eta0=0.5),
    "softmax_bf16": dict(s=1, k=6, rows=8 * 256, d=96, model="softmax", classes=5, eta0=0.5, dtype="bf16"),
    # multi-column model on one-hot features (csrc/grad_ell_mc.cu) and dense workers that abandon decided rounds
    "sparse_softmax": dict(s=1, k=6, rows=8 * 150, d=700, model="softmax", classes=5, eta0=0.5),
    "abandon": dict(s=1, k=5, abandon=True),
}


def spaced_delays(rounds, n, seed=0, gap=3e-3):
    rng = np.random.RandomState(seed)
    T = np.zeros((rounds, n))
    for r in range(rounds):
        T[r, rng.permutation(n)] = gap * (1 + np.arange(n))
    return T


CURRENT = {"eng": None, "name": ""}


def _hang_watch(secs):
    """EH_DEBUG_HANG=secs: after that long, dump the Python stacks and the flag words of the running engine."""
    import faulthandler
    import threading
    import time
    faulthandler.dump_traceback_later(secs, exit=False)

    def watch():
        time.sleep(secs)
        eng = CURRENT["eng"]
        br = getattr(eng, "bridge", None)
        if br is not None:
            print(f"[hang-watch] scheme={CURRENT['name']} {br.debug_state()}", flush=True)
    threading.Thread(target=watch, daemon=True).start()


def main():
    device = sys.argv[1]
    ctx = init_distributed(device)
    if os.environ.get("EH_DEBUG_HANG"):
        _hang_watch(float(os.environ["EH_DEBUG_HANG"]))
    bad = 0
    for name in sys.argv[2:]:
        kw = dict(KW[name])
        scheme = "approx" if name in ("sparse", "softmax", "softmax_bf16", "sparse_softmax", "abandon") else name
        n, s, P, k = kw.pop("n", 8), kw.pop("s", 1), kw.pop("P", 0), kw.pop("k", 0)
        rows, d = kw.pop("rows", 1920), kw.pop("d", 48)
        ic, pp, cv = triple_from_scheme(scheme, P)
        cfg = RunConfig(n + 1, rows, d, "", 0, "x", ic, s, pp, cv, k, 1, "AGD", rounds=6,
                        gpus=(-1 if device == "cpu" else 0),
                        beta_push=os.environ.get("EH_TEST_BETA_PUSH", "p2p"), **kw)
        if name == "sparse":
            cfg.is_real = 1
            data = generate_onehot(rows, d, 13, 8, seed=2)
        elif name == "sparse_softmax":
            cfg.is_real = 1
            data = generate_onehot(rows, d, 13, 8, seed=2, task="softmax", classes=5)
        elif name.startswith("softmax"):
            data = generate_gmm(rows, d, cfg.n_partitions, seed=1, task="softmax", classes=5)
            if name == "softmax_bf16":                   # the oracle sees the same bf16-rounded features
                import torch
                data.parts = [torch.from_numpy(q).to(torch.bfloat16).double().numpy() for q in data.parts]
        else:
            data = generate_gmm(rows, d, cfg.n_partitions, seed=1)
        dl = spaced_delays(cfg.rounds, n, seed=5)
        eng = Engine(cfg, data, ctx=ctx, delays=dl)
        CURRENT["eng"], CURRENT["name"] = eng, name
        res = eng.run()
        if ctx.is_root:
            ref = oracle.run_rounds(cfg, build_scheme(cfg, eng.spec.B), data, dl, beta0=eng.beta0, model=eng.model)
            err = np.abs(res.betaset - ref.betaset).max() / max(1.0, np.abs(ref.betaset).max())
            ok = (err <= ({"softmax": 5e-3, "softmax_bf16": 5e-3, "sparse_softmax": 5e-3}.get(name, 1e-3)) and np.array_equal(res.n_collected, ref.n_collected)
                  and np.array_equal(res.worker_timeset >= 0, ref.worker_timeset >= 0))
            print(f"[mp_engine_check] {name}: world={ctx.world} device={ctx.device} rel_err={err:.2e} "
                  f"round_ms={res.timeset.mean() * 1e3:.2f} {'OK' if ok else 'MISMATCH'}", flush=True)
            bad += 0 if ok else 1
        CURRENT["eng"] = None
        eng.close()
        ctx.barrier()
    if ctx.world > 1:
        import torch.distributed as dist
        flag = [bad]
        dist.broadcast_object_list(flag, src=0)
        bad = flag[0]
        dist.destroy_process_group()
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
