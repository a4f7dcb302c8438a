"""Scheme descriptors vs the reference semantics table."""
import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from erasurehead_b200 import RunConfig, build_scheme, scheme_from_triple, triple_from_scheme
from erasurehead_b200.config import parse_argv
from erasurehead_b200.schemes import (DECODE_FIRST_PER_GROUP, DECODE_LSTSQ, DECODE_SUM, INF, MSG_CODED,
                                      MSG_PRIVATE)


def mk(scheme, n=8, s=1, P=0, k=0, rows=960, **kw):
    ic, pp, cv = triple_from_scheme(scheme, P)
    return RunConfig(n + 1, rows, 8, "", 0, "x", ic, s, pp, cv, k, 1, "AGD", **kw)


def test_dispatch_triple_roundtrip():
    # main.py:62-92
    assert scheme_from_triple(0, 5, 3) == "naive"
    assert scheme_from_triple(1, 0, 0) == "coded"
    assert scheme_from_triple(1, 0, 1) == "replication"
    assert scheme_from_triple(1, 0, 2) == "avoidstragg"
    assert scheme_from_triple(1, 0, 3) == "approx"
    assert scheme_from_triple(1, 4, 0) == "partial_coded"
    assert scheme_from_triple(1, 4, 1) == "partial_replication"
    for sch in ("naive", "coded", "replication", "avoidstragg", "approx"):
        assert scheme_from_triple(*triple_from_scheme(sch)) == sch


def test_cli_13_positionals_and_overrides():
    cfg = parse_argv("9 8000 64 /data 0 whatever 1 1 0 3 6 1 AGD --rounds=7 --lr-schedule=expdecay --compat".split())
    assert (cfg.n_procs, cfg.n_rows, cfg.n_cols, cfg.num_collect, cfg.add_delay) == (9, 8000, 64, 6, 1)
    assert cfg.scheme == "approx" and cfg.rounds == 7 and cfg.compat
    assert cfg.dataset == "artificial-data/8000x64"                 # main.py:59-60
    assert cfg.data_dir == "/data/artificial-data/8000x64/8/"       # main.py:72
    assert cfg.alpha == pytest.approx(1 / 8000)                     # main.py:34
    assert np.allclose(cfg.learning_rates(), [10 * 0.98 ** i for i in range(1, 8)])
    with pytest.raises(SystemExit):
        parse_argv(["9", "10"])


def test_partial_layout_dir_and_linear_switch():
    cfg = parse_argv("9 8000 64 /d 1 kc_house_data 1 1 4 0 0 0 AGD".split())
    assert cfg.scheme == "partial_coded" and cfg.data_dir == "/d/kc_house_data/partial/24/"   # main.py:66
    assert cfg.model_name == "logistic"              # linear only for naive/replication/approx
    cfg = parse_argv("9 8000 64 /d 1 kc_house_data 1 1 0 3 4 0 GD".split())
    assert cfg.model_name == "linear"                # main.py:83


def test_naive_and_avoidstragg():
    sp = build_scheme(mk("naive", s=0))
    assert [sg.partition for sg in sp.segments[3]] == [3] and sp.need_count == 8 and sp.decode == DECODE_SUM
    assert sp.beta0 == "randn" and sp.prefix == "naive_acc"
    sp = build_scheme(mk("avoidstragg", s=2))
    assert sp.need_count == 6 and sp.lr_scale == pytest.approx(8 / 6) and sp.beta0 == "zeros"
    assert sp.prefix == "avoidstragg_acc_2"


def test_cyclic_assignment_and_coefficients():
    sp = build_scheme(mk("coded", s=2, seed=4))
    for w in range(8):
        assert [sg.partition for sg in sp.segments[w]] == [(w + i) % 8 for i in range(3)]   # coded.py:33-35
        for sg in sp.segments[w]:
            assert sg.coeff == sp.B[w, sg.partition] and sg.msg == MSG_CODED               # coded.py:95
    assert sp.need_count == 6 and sp.decode == DECODE_LSTSQ and sp.redundancy() == 3.0


def test_frc_assignment():
    sp = build_scheme(mk("replication", s=1))
    # replication.py:46-52: group a, rotated by position b
    assert [sg.partition for sg in sp.segments[4]] == [4, 5]
    assert [sg.partition for sg in sp.segments[5]] == [5, 4]
    assert sp.n_groups == 4 and sp.need_groups == 4 and sp.need_count == INF
    assert sp.decode == DECODE_FIRST_PER_GROUP and list(sp.group_of) == [0, 0, 1, 1, 2, 2, 3, 3]
    with pytest.raises(ValueError):
        build_scheme(mk("replication", n=7, s=1))                                          # replication.py:24-26
    sp = build_scheme(mk("approx", s=1, k=5))
    assert sp.need_count == 5 and sp.need_groups == 4 and sp.prefix == "approx_acc_1_5"
    assert build_scheme(mk("approx", s=1, k=5, compat=True)).prefix == "replication_acc_1"  # Q2


def test_partial_layouts():
    n, s, P = 4, 1, 4
    sp = build_scheme(mk("partial_coded", n=n, s=s, P=P, rows=12 * 10, seed=2))
    assert sp.n_partitions == (P - s) * n and sp.n_msgs == 2 and sp.rows_per_part == 10
    n_sep = P - s - 1
    for w in range(n):
        priv = [sg.partition for sg in sp.segments[w] if sg.msg == MSG_PRIVATE]
        cod = [sg.partition for sg in sp.segments[w] if sg.msg == MSG_CODED]
        assert priv == [i + n_sep * w for i in range(n_sep)]                               # partial_coded.py:33-36
        assert cod == [n_sep * n + (w + j) % n for j in range(s + 1)]                      # :38-43
    assert sp.need_private == n and sp.need_count == n - s
    sp = build_scheme(mk("partial_replication", n=n, s=s, P=P, rows=120))
    assert [sg.partition for sg in sp.segments[3] if sg.msg == MSG_CODED] == [8 + 2, 8 + 3]  # :44-50
    assert sp.need_private == n and sp.need_groups == 2


@settings(max_examples=60, deadline=None)
@given(st.sampled_from(["naive", "coded", "replication", "avoidstragg", "approx", "partial_coded",
                        "partial_replication"]),
       st.integers(1, 5), st.integers(1, 4), st.integers(0, 3))
def test_every_partition_is_covered(scheme, groups, sp1, extra):
    """Property: the union of all workers' partitions is the whole dataset, each coded
    partition is held by exactly s+1 workers and each private partition by exactly one."""
    s = sp1 - 1 if scheme != "naive" else 0
    n = groups * (s + 1)
    if s >= n:
        return
    P = s + 1 + extra if scheme.startswith("partial") else 0
    cfg = mk(scheme, n=n, s=s, P=P, k=max(1, n // 2), rows=n * max(1, P - s if P else 1) * 4, seed=1)
    spc = build_scheme(cfg)
    count = np.zeros(spc.n_partitions, int)
    for w in range(n):
        for sg in spc.segments[w]:
            count[sg.partition] += 1
    n_priv = sum(1 for w in range(n) for sg in spc.segments[w] if sg.msg == MSG_PRIVATE)
    assert np.all(count[:n_priv] == 1)
    red = 1 if scheme in ("naive", "avoidstragg") else s + 1
    assert np.all(count[n_priv:] == red)
    assert spc.stop(spc.need_private, n, spc.n_groups)          # everything arrived -> must stop
    assert not spc.stop(0, 0, 0) or n == 0
