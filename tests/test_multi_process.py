"""World-size > 1: gloo on CPU for the host-side logic, NCCL + peer-mapped memory on GPUs."""
import os
import socket
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ALL = ["naive", "coded", "replication", "avoidstragg", "approx", "partial_coded", "partial_replication", "sparse",
       "softmax", "softmax_bf16", "sparse_softmax", "abandon"]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _torchrun(nproc, device, schemes, timeout=300, extra_env=None):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(HERE, "mp_engine_check.py"), device] + schemes
    env = dict(os.environ, OMP_NUM_THREADS="1", **(extra_env or {}))
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env)
    return p.returncode, p.stdout + p.stderr


def test_two_process_gloo_cpu():
    rc, out = _torchrun(2, "cpu", ["approx", "coded", "partial_coded", "softmax", "sparse_softmax"])
    assert rc == 0, out[-3000:]
    assert out.count("OK") == 5


def _gpu_count():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.gpu
@pytest.mark.skipif(_gpu_count() < 2, reason="needs >= 2 GPUs")
def test_two_gpu_peer_path_all_schemes():
    rc, out = _torchrun(2, "cuda", ALL)
    assert rc == 0, out[-3000:]
    assert out.count("OK") == len(ALL)


@pytest.mark.gpu
@pytest.mark.skipif(_gpu_count() < 8, reason="needs 8 GPUs")
def test_eight_gpu_peer_path_all_schemes():
    rc, out = _torchrun(8, "cuda", ALL)
    assert rc == 0, out[-3000:]
    assert out.count("OK") == len(ALL)


@pytest.mark.gpu
@pytest.mark.skipif(_gpu_count() < 2, reason="needs >= 2 GPUs")
def test_two_gpu_multicast_parameter_push():
    """NVLS path: the decoder pushes parameters with multimem.st into a multicast-mapped buffer."""
    rc, out = _torchrun(2, "cuda", ["approx", "coded", "partial_coded"], extra_env={"EH_TEST_BETA_PUSH": "multicast"})
    assert rc == 0, out[-3000:]
    assert out.count("OK") == 3


BRIDGED = ["naive", "coded", "replication", "approx", "partial_coded", "partial_replication", "sparse", "softmax",
           "sparse_softmax"]


@pytest.mark.gpu
@pytest.mark.skipif(_gpu_count() < 2, reason="needs >= 2 GPUs")
def test_two_gpu_inter_node_bridge():
    """EH_NODE_SIZE=1 declares every GPU its own node: rank 1 is reached through the NCCL bridge
    (parallel/bridge.py) instead of the peer-mapped workspace; iterates and arrival sets must not change."""
    rc, out = _torchrun(2, "cuda", BRIDGED, extra_env={"EH_NODE_SIZE": "1"})
    assert rc == 0, out[-3000:]
    assert out.count("OK") == len(BRIDGED)


@pytest.mark.gpu
@pytest.mark.skipif(_gpu_count() < 4, reason="needs >= 4 GPUs")
def test_four_gpu_two_nodes_of_two():
    """Two GPUs in the root's NVLink domain (peer stores) + two remote GPUs (bridge) in one job."""
    rc, out = _torchrun(4, "cuda", BRIDGED, extra_env={"EH_NODE_SIZE": "2"})
    assert rc == 0, out[-3000:]
    assert out.count("OK") == len(BRIDGED)
