"""Golden-model behaviour: decode exactness, stop rules, update rules, delay model."""
import numpy as np
import pytest

from erasurehead_b200 import Engine, RunConfig, build_scheme, triple_from_scheme
from erasurehead_b200 import oracle
from erasurehead_b200.data import generate_gmm
from erasurehead_b200.delays import delay_table, expected_kth_wait
from erasurehead_b200.models import get_model


def mk(scheme, n=8, s=1, P=0, k=0, rows=960, d=12, rounds=12, **kw):
    ic, pp, cv = triple_from_scheme(scheme, P)
    return RunConfig(n + 1, rows, d, "", 0, "x", ic, s, pp, cv, k, 1, kw.pop("rule", "AGD"), rounds=rounds,
                     gpus=-1, **kw)


def run(cfg, data=None, **kw):
    data = data or generate_gmm(cfg.n_rows, cfg.n_cols, cfg.n_partitions, seed=1)
    eng = Engine(cfg, data, device="cpu", **kw)
    return eng, eng.run(), data


def test_delay_table_is_the_reference_stream():
    T = delay_table(5, 8)
    for i in range(5):
        np.random.seed(i)                                   # naive.py:146-147
        assert np.array_equal(T[i], np.random.exponential(0.5, 8))
    assert np.all(delay_table(3, 4, enabled=False) == 0)
    F = delay_table(4, 8, mode="fixed", n_stragglers=2, fixed_sleep=1.5)
    assert np.all((F == 1.5).sum(axis=1) == 2)
    D = delay_table(2, 4, mode="dead", dead=[1])
    assert np.all(np.isinf(D[:, 1])) and np.all(np.isfinite(D[:, 0]))
    assert expected_kth_wait(8, 8) == pytest.approx(1.359, abs=1e-3)
    assert expected_kth_wait(30, 15) == pytest.approx(0.338, abs=1e-3)


def test_update_rules_match_reference_formulas():
    rng = np.random.RandomState(0)
    b, u, g = rng.randn(5), rng.randn(5), rng.randn(5)
    eta, alpha, N = 10.0, 1e-3, 1000
    gm = eta / N
    assert np.allclose(oracle.gd_update(b, g, eta, alpha, gm), (1 - 2 * alpha * eta) * b - gm * g)   # naive.py:115
    i = 3
    theta = 2.0 / (i + 2.0)                                                                         # naive.py:118-122
    yt = (1 - theta) * b + theta * u
    bt = yt - gm * g - (2 * alpha * eta) * b
    ut = b + (bt - b) * (1 / theta)
    b2, u2 = oracle.agd_update(b, u, g, eta, alpha, gm, i)
    assert np.allclose(b2, bt) and np.allclose(u2, ut)


@pytest.mark.parametrize("rule", ["GD", "AGD"])
def test_exact_schemes_reproduce_uncoded_gd(rule):
    """coded / replication decode the *full* gradient from a subset of workers, so with the same
    beta_0 they must follow the uncoded trajectory (the point of exact gradient coding)."""
    base = mk("naive", s=0, rule=rule)
    data = generate_gmm(base.n_rows, base.n_cols, 8, seed=1)
    b0 = np.zeros(base.n_cols)
    ref = oracle.run_rounds(base, build_scheme(base), data, delay_table(base.rounds, 8), beta0=b0)
    for scheme, s in (("coded", 2), ("replication", 1), ("coded", 1)):
        cfg = mk(scheme, s=s, rule=rule, seed=3)
        res = oracle.run_rounds(cfg, build_scheme(cfg), data, delay_table(cfg.rounds, 8), beta0=b0)
        assert np.abs(res.betaset - ref.betaset).max() < 1e-6 * max(1, np.abs(ref.betaset).max()), scheme
        assert np.all(res.timeset <= ref.timeset + 1e-12)      # never slower than waiting for everyone


def test_partial_schemes_are_exact_too():
    n, s, P = 4, 1, 4
    cfg_pc = mk("partial_coded", n=n, s=s, P=P, rows=12 * 20, seed=2)
    data = generate_gmm(cfg_pc.n_rows, cfg_pc.n_cols, cfg_pc.n_partitions, seed=5)
    dl = delay_table(cfg_pc.rounds, n)
    a = oracle.run_rounds(cfg_pc, build_scheme(cfg_pc), data, dl)
    cfg_pr = mk("partial_replication", n=n, s=s, P=P, rows=12 * 20)
    b = oracle.run_rounds(cfg_pr, build_scheme(cfg_pr), data, dl)
    # both see the full gradient of the same 12 partitions
    full = mk("naive", n=12, s=0, rows=12 * 20)
    c = oracle.run_rounds(full, build_scheme(full), data, np.zeros((full.rounds, 12)), beta0=np.zeros(cfg_pc.n_cols))
    assert np.abs(a.betaset - c.betaset).max() < 1e-6
    assert np.abs(b.betaset - c.betaset).max() < 1e-9
    # stop rule: all private parts => the round lasts as long as the slowest worker
    assert np.allclose(a.timeset, dl.max(axis=1))


def test_stop_rules_and_worker_timeset():
    cfg = mk("approx", s=1, k=5)
    _, res, _ = run(cfg)
    dl = delay_table(cfg.rounds, 8)
    for r in range(cfg.rounds):
        order = np.argsort(dl[r])
        # replay approximate_coding.py:144-158
        cov, cnt, used = set(), 0, []
        for w in order:
            if cnt >= 5 or len(cov) >= 4:
                break
            cnt += 1
            if w // 2 not in cov:
                cov.add(w // 2)
                used.append(w)
        assert res.n_collected[r] == cnt
        assert sorted(np.flatnonzero(res.weights[r])) == sorted(used)
        assert res.timeset[r] == pytest.approx(dl[r][order[cnt - 1]])
        collected = set(order[:cnt])
        for w in range(8):
            if w in collected:
                assert res.worker_timeset[r, w] == pytest.approx(dl[r, w])
            else:
                assert res.worker_timeset[r, w] == -1                      # coded.py:171-173


def test_avoidstragg_rescales_and_waits_for_n_minus_s():
    cfg = mk("avoidstragg", s=2)
    _, res, _ = run(cfg)
    dl = delay_table(cfg.rounds, 8)
    assert np.allclose(res.timeset, np.sort(dl, axis=1)[:, 5])
    assert np.all(res.n_collected == 6)


def test_mean_wait_matches_order_statistics():
    cfg = mk("coded", s=2, rounds=400, rows=96, d=3, seed=1)
    _, res, _ = run(cfg)
    assert res.timeset.mean() == pytest.approx(expected_kth_wait(8, 6), rel=0.1)


def test_dead_worker_tolerated_by_coded_but_not_by_naive():
    cfg = mk("coded", s=2, seed=1)
    dl = delay_table(cfg.rounds, 8, mode="dead", dead=[0, 5])
    data = generate_gmm(cfg.n_rows, cfg.n_cols, 8, seed=1)
    res = oracle.run_rounds(cfg, build_scheme(cfg), data, dl)
    assert np.all(res.worker_timeset[:, [0, 5]] == -1)
    base = mk("naive", s=0)
    with pytest.raises(RuntimeError):
        oracle.run_rounds(base, build_scheme(base), data, dl)


def test_linear_and_softmax_models_learn():
    cfg = mk("replication", s=1, model="linear", eta0=0.05, rounds=40)
    data = generate_gmm(cfg.n_rows, cfg.n_cols, 8, seed=1, task="linear")
    eng = Engine(cfg, data, device="cpu")
    ev = eng.evaluate(eng.run())
    assert ev.auc is None and ev.training_loss[-1] < 0.2 * ev.training_loss[0]
    cfg = mk("approx", s=1, k=6, model="softmax", classes=3, eta0=0.5, rounds=30)
    data = generate_gmm(cfg.n_rows, cfg.n_cols, 8, seed=1, task="softmax", classes=3)
    eng = Engine(cfg, data, device="cpu")
    res = eng.run()
    assert res.betaset.shape == (30, cfg.n_cols, 3)
    ev = eng.evaluate(res)
    assert ev.training_loss[-1] < ev.training_loss[0]


def test_gradient_encoding_folds_into_residual():
    rng = np.random.RandomState(0)
    X, y, b = rng.randn(20, 4), rng.choice([-1.0, 1.0], 20), rng.randn(4)
    m = get_model("logistic")
    # coded.py:184: X^T ( (c*y) / (exp(p*y)+1) ) * -1
    c = 2.5
    want = -X.T.dot((c * y) / (np.exp(X.dot(b) * y) + 1))
    assert np.allclose(m.gradient(X, y, np.full(20, c), b), want)
    ml = get_model("linear")
    assert np.allclose(ml.gradient(X, y, np.ones(20), b), -2 * X.T.dot(y - X.dot(b)))      # naive.py:345-346


def test_carryover_delays_replay_a_backlog():
    """A worker still inside its delay when the round ends starts the next round late by what is left
    (the reference's coded driver: its master never waits for stragglers, coded.py:137-166)."""
    from erasurehead_b200 import RunConfig, build_scheme, oracle, triple_from_scheme
    ic, pp, cv = triple_from_scheme("coded", 0)
    spec = build_scheme(RunConfig(4, 3 * 40, 6, "", 0, "x", ic, 1, pp, cv, 0, 1, "AGD", seed=1))      # 3 workers, waits for 2
    dl = np.array([[0.1, 0.2, 1.0],
                   [0.5, 0.5, 0.1],
                   [0.3, 0.05, 0.05]])
    eff = oracle.carryover_delays(spec, dl)
    assert np.allclose(eff[0], [0.1, 0.2, 1.0])                 # round 0 ends at 0.2, worker 2 has 0.8 s left
    assert np.allclose(eff[1], [0.5, 0.5, 0.9])                 # ... ends at 0.5, worker 2 has 0.4 s left
    assert np.allclose(eff[2], [0.3, 0.05, 0.45])
    # a scheme that waits for everybody never builds a backlog
    ic, pp, cv = triple_from_scheme("naive", 0)
    naive = build_scheme(RunConfig(4, 3 * 40, 6, "", 0, "x", ic, 0, pp, cv, 0, 1, "AGD", seed=1))
    assert np.allclose(oracle.carryover_delays(naive, dl), dl)
    # a dead worker stays dead
    dl[0, 2] = np.inf
    assert np.isinf(oracle.carryover_delays(spec, dl)[:, 2]).all()


def test_carryover_is_on_only_for_compat_coded():
    from erasurehead_b200 import RunConfig, triple_from_scheme
    mk = lambda scheme, **kw: RunConfig(5, 480, 6, "", 0, "x", *triple_from_scheme(scheme, 0)[:1], 1,
                                        *triple_from_scheme(scheme, 0)[1:], 0, 1, "AGD", **kw)
    assert mk("coded", compat=True).delay_carryover is True
    assert mk("coded").delay_carryover is False
    assert mk("replication", compat=True).delay_carryover is False
    assert mk("replication", delay_carryover=True).delay_carryover is True


def test_carryover_properties():
    """Over random tables: a backlog only ever adds delay, never for a scheme that waits for every worker, and a
    round's effective delay is the raw one exactly for the workers that were collected in the round before."""
    from hypothesis import given, settings, strategies as st
    from erasurehead_b200 import RunConfig, build_scheme, oracle, triple_from_scheme

    def spec_of(scheme, n, s, k=0):
        ic, pp, cv = triple_from_scheme(scheme, 0)
        return build_scheme(RunConfig(n + 1, n * 16, 4, "", 0, "x", ic, s, pp, cv, k, 1, "AGD", seed=1))

    specs = [spec_of("coded", 6, 2), spec_of("approx", 6, 1, 3), spec_of("avoidstragg", 6, 2), spec_of("replication", 6, 2)]

    @settings(max_examples=40, deadline=None)
    @given(st.integers(0, 2 ** 31 - 1), st.integers(0, 3))
    def check(seed, which):
        spec = specs[which]
        dl = np.random.RandomState(seed).exponential(0.5, (12, 6))
        eff = oracle.carryover_delays(spec, dl)
        assert (eff >= dl - 1e-12).all()
        for i in range(1, 12):
            wt = oracle.plan_round(spec, eff[i - 1])[4]
            t_stop = oracle.plan_round(spec, eff[i - 1])[2]
            on_time = eff[i - 1] <= t_stop + 1e-12
            assert np.allclose(eff[i][on_time], dl[i][on_time])
            assert (wt[~on_time] < 0).all()                    # whoever was still busy had not been collected
    check()
    naive = spec_of("naive", 6, 0)
    dl = np.random.RandomState(0).exponential(0.5, (12, 6))
    assert np.allclose(oracle.carryover_delays(naive, dl), dl)
