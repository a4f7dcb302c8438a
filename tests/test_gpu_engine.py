"""Full fused path on one GPU (all workers on it) against the float64 oracle."""
import numpy as np
import pytest
import torch

from erasurehead_b200 import Engine, RunConfig, build_scheme, oracle, triple_from_scheme
from erasurehead_b200.data import generate_gmm, generate_onehot

pytestmark = pytest.mark.gpu


def spaced_delays(rounds, n, seed=0, gap=2e-3):
    """Per-round random arrival order with >= `gap` seconds between arrivals, so the device's
    arrival order is exactly the table's (compute jitter is microseconds)."""
    rng = np.random.RandomState(seed)
    T = np.zeros((rounds, n))
    for r in range(rounds):
        T[r, rng.permutation(n)] = gap * (1 + np.arange(n))
    return T


def mk(scheme, n=8, s=1, P=0, k=0, rows=1920, d=48, rounds=6, **kw):
    ic, pp, cv = triple_from_scheme(scheme, P)
    return RunConfig(n + 1, rows, d, "", 0, "x", ic, s, pp, cv, k, 1, kw.pop("rule", "AGD"), rounds=rounds, **kw)


CASES = [
    ("naive", dict(s=0)), ("coded", dict(s=2, seed=3)), ("replication", dict(s=1)),
    ("avoidstragg", dict(s=2)), ("approx", dict(s=1, k=5)),
    ("partial_coded", dict(n=4, s=1, P=4, rows=12 * 160, seed=2)),
    ("partial_replication", dict(n=4, s=1, P=4, rows=12 * 160)),
]


@pytest.mark.parametrize("scheme,kw", CASES)
@pytest.mark.parametrize("rule", ["GD", "AGD"])
def test_engine_matches_oracle(scheme, kw, rule):
    cfg = mk(scheme, rule=rule, **kw)
    data = generate_gmm(cfg.n_rows, cfg.n_cols, cfg.n_partitions, seed=1)
    dl = spaced_delays(cfg.rounds, cfg.n_workers, seed=5)
    eng = Engine(cfg, data, delays=dl)
    assert eng.on_gpu
    res = eng.run()
    spec = build_scheme(cfg, eng.spec.B)
    ref = oracle.run_rounds(cfg, spec, data, dl, beta0=eng.beta0)
    scale = max(1.0, np.abs(ref.betaset).max())
    assert np.abs(res.betaset - ref.betaset).max() <= 5e-4 * scale
    assert np.array_equal(res.n_collected, ref.n_collected)
    assert np.array_equal(res.worker_timeset >= 0, ref.worker_timeset >= 0)
    assert np.allclose(res.weights, ref.weights, atol=1e-6 * max(1, np.abs(ref.weights).max()))
    # device timers: round time ~ injected delay of the last consumed message
    assert np.all(res.timeset >= ref.timeset - 1e-4) and np.all(res.timeset <= ref.timeset + 5e-3)
    ev = eng.evaluate(res)
    from erasurehead_b200 import results
    evh = results.evaluate_host(eng.model, data, res.betaset)
    # default evaluation runs on the tensor cores with TF32 operands
    assert np.allclose(ev.training_loss, evh.training_loss, rtol=2e-3, atol=1e-4)
    assert np.allclose(ev.testing_loss, evh.testing_loss, rtol=2e-3, atol=1e-4)
    assert np.allclose(ev.auc, evh.auc, atol=3e-3)
    assert np.allclose(ev.accuracy, evh.accuracy, atol=5e-3)
    eng.cfg.eval_precision = "fp32"
    ev32 = eng.evaluate(res)
    assert np.allclose(ev32.training_loss, evh.training_loss, rtol=2e-4, atol=1e-5)
    assert np.allclose(ev32.auc, evh.auc, atol=2e-3)
    eng.close()


def test_engine_solve_decode_equals_table():
    cfg = mk("coded", s=2, seed=3, decode="solve")
    data = generate_gmm(cfg.n_rows, cfg.n_cols, 8, seed=1)
    dl = spaced_delays(cfg.rounds, 8, seed=9)
    a = Engine(cfg, data, delays=dl)
    ra = a.run()
    b = Engine(cfg.replace(decode="table"), data, delays=dl)
    rb = b.run()
    assert np.allclose(ra.weights, rb.weights, rtol=1e-7, atol=1e-9)
    assert np.allclose(ra.betaset, rb.betaset, atol=1e-9)
    a.close(); b.close()


def test_engine_sparse_linear_and_logistic():
    for task, model in (("logistic", None), ("linear", "linear")):
        cfg = mk("approx", s=1, k=6, rows=8 * 120, d=700, model=model, eta0=(0.05 if model else 10.0))
        cfg.is_real, cfg.dataset = 1, "amazon-dataset"
        data = generate_onehot(cfg.n_rows, cfg.n_cols, 13, 8, seed=2, task=task)
        dl = spaced_delays(cfg.rounds, 8, seed=1)
        eng = Engine(cfg, data, delays=dl)
        res = eng.run()
        ref = oracle.run_rounds(cfg, build_scheme(cfg), data, dl, beta0=eng.beta0, model=eng.model)
        assert np.abs(res.betaset - ref.betaset).max() <= 1e-3 * max(1.0, np.abs(ref.betaset).max())
        eng.close()


def test_dead_worker_is_tolerated_and_watchdog_fires():
    cfg = mk("coded", s=2, seed=3)
    data = generate_gmm(cfg.n_rows, cfg.n_cols, 8, seed=1)
    dl = spaced_delays(cfg.rounds, 8, seed=2)
    dl[:, [1, 6]] = np.inf                               # two permanently dead workers, s = 2
    eng = Engine(cfg, data, delays=dl)
    res = eng.run()
    assert np.all(res.worker_timeset[:, [1, 6]] == -1)
    ref = oracle.run_rounds(cfg, build_scheme(cfg, eng.spec.B), data, dl, beta0=eng.beta0)
    assert np.abs(res.betaset - ref.betaset).max() <= 5e-4 * max(1.0, np.abs(ref.betaset).max())
    eng.close()
    # naive cannot finish a round with a dead worker: the device watchdog must fire, not hang
    cfg = mk("naive", s=0, rounds=2)
    dl = np.zeros((2, 8)); dl[:, 3] = np.inf
    eng = Engine(cfg, data, delays=dl)
    eng.timeout_ns = eng.dec.timeout_ns = eng.plan.timeout_ns = int(0.3e9)
    with pytest.raises(RuntimeError, match="watchdog"):
        eng.run()
    eng.close()


def test_synthetic_on_device_and_checkpoint_resume(tmp_path):
    base = dict(s=1, k=5, rows=8 * 2048, d=256, synthetic=True, seed=4, output_dir=str(tmp_path) + "/")
    cfg = mk("approx", rounds=8, **base)
    dl = spaced_delays(8, 8, seed=3)
    full = Engine(cfg, None, delays=dl)
    rf = full.run()
    ev = full.evaluate(rf)
    assert ev.training_loss[-1] < ev.training_loss[0] and ev.auc[-1] > 0.8
    full.close()
    first = Engine(mk("approx", rounds=4, **base), None, delays=dl[:4])
    first.run()
    ck = str(tmp_path / "ck.npz")
    first.save_checkpoint(ck, 4)
    first.close()
    second = Engine(mk("approx", rounds=8, resume=ck, **base), None, delays=dl[4:])
    r2 = second.run()
    assert np.allclose(r2.betaset, rf.betaset[4:], atol=1e-6)
    second.close()


def test_nccl_baseline_engine_matches_oracle_single_gpu():
    """Baseline B2 (torch + cuBLAS, library calls only) follows the same trajectory."""
    from erasurehead_b200.parallel.nccl_baseline import NcclBaselineEngine
    cfg = mk("coded", s=2, seed=3, rounds=4)
    data = generate_gmm(cfg.n_rows, cfg.n_cols, 8, seed=1)
    dl = spaced_delays(cfg.rounds, 8, seed=5, gap=1e-4)
    eng = NcclBaselineEngine(cfg, data, delays=dl)
    eng.prepare(); eng.enqueue(4); eng.finish()
    ref = oracle.run_rounds(cfg, build_scheme(cfg), data, dl)
    assert np.abs(eng.beta.cpu().numpy() - ref.betaset[-1]).max() <= 5e-4 * max(1.0, np.abs(ref.betaset).max())


def test_engine_softmax_multiclass_matches_oracle():
    """Multi-class softmax regression through the full coded path (tcgen05 gradient kernels)."""
    C_ = 5
    cfg = mk("approx", s=1, k=6, rows=8 * 512, d=160, model="softmax", classes=C_, eta0=0.5, rounds=5)
    data = generate_gmm(cfg.n_rows, cfg.n_cols, 8, seed=1, task="softmax", classes=C_)
    dl = spaced_delays(cfg.rounds, 8, seed=4)
    eng = Engine(cfg, data, delays=dl)
    res = eng.run()
    assert res.betaset.shape == (5, 160, C_)
    ref = oracle.run_rounds(cfg, build_scheme(cfg), data, dl, beta0=eng.beta0, model=eng.model)
    scale = max(1.0, np.abs(ref.betaset).max())
    assert np.abs(res.betaset - ref.betaset).max() <= 5e-3 * scale        # TF32 operands
    assert np.array_equal(res.n_collected, ref.n_collected)
    ev = eng.evaluate(res)
    from erasurehead_b200 import results
    evh = results.evaluate_host(eng.model, data, res.betaset)
    assert np.allclose(ev.training_loss, evh.training_loss, rtol=3e-3)
    assert np.allclose(ev.testing_loss, evh.testing_loss, rtol=3e-3)
    assert np.allclose(ev.accuracy, evh.accuracy, atol=5e-3) and ev.accuracy[-1] > 1.5 / C_
    assert ev.training_loss[-1] < ev.training_loss[0]
    eng.close()
    # exact coded scheme on the same data: decode weights act on the flat multi-column vector
    cfg2 = mk("coded", s=2, seed=3, rows=8 * 512, d=160, model="softmax", classes=C_, eta0=0.5, rounds=4)
    eng2 = Engine(cfg2, data, delays=spaced_delays(4, 8, seed=6))
    r2 = eng2.run()
    ref2 = oracle.run_rounds(cfg2, build_scheme(cfg2, eng2.spec.B), data, eng2.delays, beta0=eng2.beta0, model=eng2.model)
    assert np.abs(r2.betaset - ref2.betaset).max() <= 1e-2 * max(1.0, np.abs(ref2.betaset).max())
    eng2.close()


def test_engine_softmax_bf16_rows_matches_oracle_on_rounded_data():
    """--dtype=bf16 with a multi-class model: the all-tcgen05 one-pass kernel (csrc/grad_mc_bf16.cu)."""
    import torch
    C_ = 6
    cfg = mk("approx", s=1, k=6, rows=8 * 512, d=200, model="softmax", classes=C_, eta0=0.5, rounds=5, dtype="bf16")
    data = generate_gmm(cfg.n_rows, cfg.n_cols, 8, seed=1, task="softmax", classes=C_)
    data.parts = [torch.from_numpy(p).to(torch.bfloat16).double().numpy() for p in data.parts]
    dl = spaced_delays(cfg.rounds, 8, seed=4)
    eng = Engine(cfg, data, delays=dl)
    assert eng.bf16 and eng.mc and eng.parts[0][0].dtype == torch.bfloat16
    res = eng.run()
    ref = oracle.run_rounds(cfg, build_scheme(cfg), data, dl, beta0=eng.beta0, model=eng.model)
    scale = max(1.0, np.abs(ref.betaset).max())
    assert np.abs(res.betaset - ref.betaset).max() <= 5e-3 * scale        # hi + lo bf16 operand pairs for W and the residual
    assert np.array_equal(res.n_collected, ref.n_collected)
    ev = eng.evaluate(res)
    assert ev.training_loss[-1] < ev.training_loss[0]
    eng.close()


def test_debug_poison_mode_runs_clean():
    """NaN-poisoned slots: a correct run never reads a slot nobody wrote (sanitizer substitute)."""
    cfg = mk("approx", s=1, k=5, debug_poison=True)
    data = generate_gmm(cfg.n_rows, cfg.n_cols, 8, seed=1)
    eng = Engine(cfg, data, delays=spaced_delays(cfg.rounds, 8, seed=5))
    res = eng.run()
    assert np.all(np.isfinite(res.betaset))
    eng.close()


@pytest.mark.parametrize("scheme,kw", [("coded", dict(n=30, s=3, seed=7)), ("approx", dict(n=30, s=2, k=15)),
                                       ("replication", dict(n=30, s=4)), ("avoidstragg", dict(n=30, s=3)),
                                       ("partial_coded", dict(n=12, s=2, P=5, rows=36 * 40, seed=5))])
def test_reference_cluster_sizes_30_workers(scheme, kw):
    """The reference's default cluster is 30 workers (run_approx_coding.sh:2): the decode table has
    C(30,3) = 4060 rows and cond(B) reaches 1e5 -- float64 decode on the device."""
    rows = kw.pop("rows", 30 * 64)
    cfg = mk(scheme, rows=rows, d=40, rounds=5, **kw)
    data = generate_gmm(cfg.n_rows, cfg.n_cols, cfg.n_partitions, seed=2)
    dl = spaced_delays(cfg.rounds, cfg.n_workers, seed=11, gap=1e-3)
    eng = Engine(cfg, data, delays=dl)
    res = eng.run()
    ref = oracle.run_rounds(cfg, build_scheme(cfg, eng.spec.B), data, dl, beta0=eng.beta0)
    scale = max(1.0, np.abs(ref.betaset).max())
    # messages are fp32 and |decode weights| reach ~1e2-1e4 for (30,3): tolerance scales with them
    tol = 5e-4 * max(1.0, np.abs(ref.weights).max())
    assert np.abs(res.betaset - ref.betaset).max() <= tol * scale
    assert np.array_equal(res.n_collected, ref.n_collected)
    assert np.array_equal(res.worker_timeset >= 0, ref.worker_timeset >= 0)
    eng.close()


def test_random_scheme_configurations_match_oracle():
    """Randomised (n, s, k, P) sweep of the device decoder against the oracle."""
    rng = np.random.RandomState(123)
    for trial in range(10):
        scheme = ["naive", "coded", "replication", "avoidstragg", "approx", "partial_coded", "partial_replication"][trial % 7]
        s = int(rng.randint(1, 4)) if scheme != "naive" else 0
        groups = int(rng.randint(2, 5))
        n = groups * (s + 1)
        P = s + 1 + int(rng.randint(1, 3)) if scheme.startswith("partial") else 0
        k = int(rng.randint(1, n + 1))
        parts = (P - s) * n if P else n
        cfg = mk(scheme, n=n, s=s, P=P, k=k, rows=parts * 48, d=24, rounds=4, seed=trial, rule=("GD" if trial % 2 else "AGD"))
        data = generate_gmm(cfg.n_rows, cfg.n_cols, cfg.n_partitions, seed=trial)
        dl = spaced_delays(cfg.rounds, n, seed=trial, gap=1e-3)
        eng = Engine(cfg, data, delays=dl)
        res = eng.run()
        ref = oracle.run_rounds(cfg, build_scheme(cfg, eng.spec.B), data, dl, beta0=eng.beta0)
        tol = 5e-4 * max(1.0, np.abs(ref.weights).max()) * max(1.0, np.abs(ref.betaset).max())
        assert np.abs(res.betaset - ref.betaset).max() <= tol, (scheme, n, s, k, P)
        assert np.array_equal(res.n_collected, ref.n_collected), (scheme, n, s, k, P)
        eng.close()


def test_engine_bf16_storage_matches_oracle_on_rounded_data():
    cfg = mk("approx", s=1, k=5, d=1200, rows=8 * 300, dtype="bf16")
    data = generate_gmm(cfg.n_rows, cfg.n_cols, 8, seed=1)
    # the oracle sees the same bf16-rounded features
    data.parts = [torch.from_numpy(p).to(torch.bfloat16).double().numpy() for p in data.parts]
    dl = spaced_delays(cfg.rounds, 8, seed=5)
    eng = Engine(cfg, data, delays=dl)
    assert eng.bf16 and eng.parts[0][0].dtype == torch.bfloat16
    res = eng.run()
    ref = oracle.run_rounds(cfg, build_scheme(cfg), data, dl, beta0=eng.beta0)
    assert np.abs(res.betaset - ref.betaset).max() <= 5e-4 * max(1.0, np.abs(ref.betaset).max())
    ev = eng.evaluate(res)
    assert np.all(np.isfinite(ev.training_loss))
    eng.close()


@pytest.mark.parametrize("model", ["logistic", "linear"])
def test_engine_narrow_bf16_rows_take_the_tensor_core_kernel(model):
    """One worker per GPU and narrow bf16 rows: the single-column gradient runs on the all-tcgen05 kernel
    (one message per launch); same oracle agreement as the CUDA-core bf16 path."""
    cfg = mk("naive", n=1, s=0, d=200, rows=3000, dtype="bf16", model=model, eta0=0.5 if model == "linear" else 10.0)
    task = "linear" if model == "linear" else "logistic"
    data = generate_gmm(cfg.n_rows, cfg.n_cols, 1, seed=2, task=task)
    data.parts = [torch.from_numpy(p).to(torch.bfloat16).double().numpy() for p in data.parts]
    dl = spaced_delays(cfg.rounds, 1, seed=1)
    eng = Engine(cfg, data, delays=dl)
    assert eng.bf16 and eng.n_jobs == 1
    res = eng.run()
    ref = oracle.run_rounds(cfg, build_scheme(cfg), data, dl, beta0=eng.beta0, model=eng.model)
    assert np.abs(res.betaset - ref.betaset).max() <= 5e-4 * max(1.0, np.abs(ref.betaset).max())
    eng.close()


def test_cuda_graph_mode_matches_stream_mode():
    """The whole run captured into one CUDA graph gives the same trajectory."""
    for scheme, kw in (("approx", dict(s=1, k=5)), ("coded", dict(s=2, seed=3)), ("partial_coded", dict(n=4, s=1, P=4, rows=12 * 160, seed=2))):
        cfg = mk(scheme, **kw)
        data = generate_gmm(cfg.n_rows, cfg.n_cols, cfg.n_partitions, seed=1)
        dl = spaced_delays(cfg.rounds, cfg.n_workers, seed=5)
        a = Engine(cfg, data, delays=dl); ra = a.run(); a.close()
        b = Engine(cfg.replace(cuda_graph=True), data, delays=dl); rb = b.run(); b.close()
        assert np.array_equal(ra.betaset, rb.betaset) and np.array_equal(ra.n_collected, rb.n_collected)


def test_arrival_is_completion_plus_delay_and_decoder_overlaps_compute():
    """Compute >> delay: every gradient message is ~0.7 ms of HBM streaming in its own launch (4 GiB), the
    injected delays are tens of microseconds.  A message must become ready at its worker's own completion
    stamp + its delay (the reference worker computes, sleeps, then sends: approximate_coding.py:194-207),
    and the decoder -- on its own stream, on SMs the gradient kernels leave free -- must fire the stop rule
    while later workers of the same GPU are still computing (approximate_coding.py:144-158: rank 0 only waits)."""
    n, s, k, d, rpp, R = 4, 1, 2, 1024, 524288, 4
    cfg = mk("approx", n=n, s=s, k=k, rows=n * rpp, d=d, rounds=R, synthetic=True, seed=1)
    dl = np.tile(np.array([300e-6, 10e-6, 10e-6, 10e-6]), (R, 1))
    eng = Engine(cfg, None, delays=dl)
    assert eng.dec_ctas >= 1 and eng._dec_stream is not None
    res = eng.run()
    done, wts = res.done, res.worker_timeset
    job = done[0, 1] - done[0, 0]                      # one message = one launch; they run back to back
    assert 3e-4 < job < 3e-3, job
    for r in range(R):
        got = wts[r] >= 0
        assert got.sum() == k and got[0] and got[1]    # 0 and 1 finish first and cover k = 2
        # ready = completion + delay (polling granularity and clamp: a few microseconds)
        assert np.all(np.abs(wts[r, got] - (done[r, got] + dl[r, got])) < 30e-6), (wts[r], done[r])
    # round 0: the stop rule fires right after worker 1's message, long before workers 2 and 3 finish
    assert res.timeset[0] < done[0, 1] + 10e-6 + 60e-6
    assert res.timeset[0] < 2.6 * job
    eng.close()
    # same run with the decoder serialised behind the gradient kernels: the round waits for all four
    eng2 = Engine(cfg.replace(decoder="inline"), None, delays=dl)
    res2 = eng2.run()
    assert res2.timeset[0] > 3.5 * job
    assert np.allclose(res2.betaset, res.betaset, atol=1e-6)     # 147 vs 148 gradient CTAs: other summation order
    eng2.close()


@pytest.mark.parametrize("scheme,kw", [("approx", dict(s=1, k=5)), ("coded", dict(s=2, seed=3))])
def test_engine_softmax_on_onehot_features_matches_oracle(scheme, kw):
    """Multi-class model on one-hot (ELL) features -- 'Covertype-shaped multi-class' -- against the oracle."""
    cfg = mk(scheme, rows=8 * 300, d=900, model="softmax", classes=7, **kw)
    data = generate_onehot(cfg.n_rows, cfg.n_cols, 12, cfg.n_partitions, seed=4, task="softmax", classes=7)
    dl = spaced_delays(cfg.rounds, cfg.n_workers, seed=6)
    eng = Engine(cfg, data, delays=dl)
    assert eng.on_gpu and eng.sparse and eng.mc
    res = eng.run()
    ref = oracle.run_rounds(cfg, build_scheme(cfg, eng.spec.B), data, dl, beta0=eng.beta0)
    assert res.betaset.shape == ref.betaset.shape == (cfg.rounds, 900, 7)
    scale = max(1.0, np.abs(ref.betaset).max())
    assert np.abs(res.betaset - ref.betaset).max() <= 5e-4 * scale
    assert np.array_equal(res.worker_timeset >= 0, ref.worker_timeset >= 0)
    ev = eng.evaluate(res)
    from erasurehead_b200 import results
    evh = results.evaluate_host(eng.model, data, res.betaset)
    assert np.allclose(ev.training_loss, evh.training_loss, rtol=1e-4, atol=1e-5)
    assert np.allclose(ev.testing_loss, evh.testing_loss, rtol=1e-4, atol=1e-5)
    eng.close()


@pytest.mark.parametrize("model,classes", [("logistic", 1), ("softmax", 7)])
def test_engine_runs_on_device_generated_onehot_shapes(model, classes):
    """--synthetic with is_real=1: one-hot data of the Covertype shape generated on the GPU (no host data),
    binary and 7-class; the loss must go down."""
    n, s, k = 6, 2, 3
    ic, pp, cv = triple_from_scheme("approx")
    cfg = RunConfig(n + 1, 6 * 4000, 15509, "", 1, "covtype", ic, s, pp, cv, k, 0, "AGD", rounds=12, synthetic=True,
                    seed=2, model=model, classes=classes, init="zeros", eta0=2.0)
    eng = Engine(cfg, None)
    assert eng.sparse and eng.synth_shape == {"nnz": 55}
    res = eng.run()
    ev = eng.evaluate(res)
    assert np.all(np.isfinite(res.betaset))
    assert ev.training_loss[-1] < ev.training_loss[0]
    eng.close()


def test_abandon_skips_the_compute_of_workers_whose_round_is_over():
    """--abandon: a gradient kernel that is still streaming when the stop rule fires stops reading X and ships
    nothing (the reference can only cancel the send, README.md:122).  Four 4 GiB messages back to back on one
    GPU, n_collect = 2: workers 2 and 3 are abandoned every round, so a round costs about two messages
    instead of four, with the same iterates."""
    n, s, k, d, rpp, R = 4, 1, 2, 1024, 524288, 5
    cfg = mk("approx", n=n, s=s, k=k, rows=n * rpp, d=d, rounds=R, synthetic=True, seed=1)
    dl = np.tile(np.array([20e-6, 10e-6, 10e-6, 10e-6]), (R, 1))
    plain = Engine(cfg, None, delays=dl)
    r0 = plain.run()
    plain.close()
    eng = Engine(cfg.replace(abandon=True), None, delays=dl)
    r1 = eng.run()
    eng.close()
    assert np.allclose(r1.betaset, r0.betaset, atol=1e-6)
    assert np.array_equal(r1.worker_timeset >= 0, r0.worker_timeset >= 0)
    job = r1.done[1, 1] - r1.done[1, 0]
    # without abandoning, worker 2's kernel starts before the decoder has published the new round and runs to the
    # end (worker 3's then skips at launch): ~3 messages of streaming per round
    assert r0.timeset[1:].mean() > 2.6 * job
    # with it, worker 2 stops within a few tiles: ~2 messages per round
    assert r1.timeset[1:].mean() < 2.3 * job
    assert np.all(r1.done[1:, 2:] < 0)           # their messages never arrived
