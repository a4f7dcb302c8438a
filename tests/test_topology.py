"""Node topology of a job and the message order of the inter-node bridge (host-side logic, no GPU)."""
import os
import subprocess
import sys

from erasurehead_b200 import RunConfig, build_scheme, triple_from_scheme
from erasurehead_b200.parallel.bridge import messages_of
from erasurehead_b200.parallel.context import DistContext, node_ids
from erasurehead_b200.schemes import MSG_CODED, MSG_PRIVATE

HERE = os.path.dirname(os.path.abspath(__file__))


def test_single_node_has_no_remote_ranks():
    ctx = DistContext(rank=1, world=4)
    assert ctx.domain_ranks == [0, 1, 2, 3] and ctx.remote_ranks == [] and not ctx.is_remote


def test_domain_is_the_roots_node():
    node_of = [0, 0, 1, 1, 2, 2]
    for rank in range(6):
        ctx = DistContext(rank=rank, world=6, node_of=node_of)
        assert ctx.domain_ranks == [0, 1]
        assert ctx.remote_ranks == [2, 3, 4, 5]
        assert ctx.is_remote == (rank >= 2)


def test_node_size_override(monkeypatch):
    monkeypatch.setenv("EH_NODE_SIZE", "2")
    assert node_ids(0, 4) == [0, 0, 1, 1]
    assert node_ids(0, 2) is None                      # everything on one node
    assert node_ids(0, 1) is None
    monkeypatch.setenv("EH_NODE_SIZE", "1")
    assert node_ids(1, 3) == [0, 1, 2]


def _spec(scheme, n, s, P=0, k=0):
    ic, pp, cv = triple_from_scheme(scheme, P)
    rows = 12 * 16 * n
    return build_scheme(RunConfig(n + 1, rows, 8, "", 0, "x", ic, s, pp, cv, k, 0, "GD", seed=1))


def test_bridge_message_order_matches_job_order():
    """Both ends of a pair enumerate a rank's messages identically: per local worker, private part first."""
    ctx = DistContext(rank=0, world=4, node_of=[0, 0, 1, 1])
    spec = _spec("approx", 8, 1, k=5)
    assert messages_of(spec, ctx, 2) == [(2, MSG_CODED), (6, MSG_CODED)]
    assert messages_of(spec, ctx, 3) == [(3, MSG_CODED), (7, MSG_CODED)]
    spec = _spec("partial_coded", 4, 1, P=4)
    assert messages_of(spec, ctx, 3) == [(3, MSG_PRIVATE), (3, MSG_CODED)]
    # every message of the job is sent by exactly one rank
    allm = [m for r in range(4) for m in messages_of(spec, ctx, r)]
    assert sorted(allm) == sorted((w, k) for w in range(4) for k in (MSG_CODED, MSG_PRIVATE))


def test_cpu_job_ignores_node_layout():
    """The float64 CPU path has no transport: a declared multi-node layout must not change it."""
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(HERE, "mp_engine_check.py"), "cpu", "approx", "partial_coded"]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, OMP_NUM_THREADS="1", EH_NODE_SIZE="1"))
    assert p.returncode == 0, (p.stdout + p.stderr)[-3000:]
    assert p.stdout.count("OK") == 2
