"""Cluster tool: argument plumbing and host-file outputs (no network)."""
import os
import subprocess
import sys

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def test_get_hosts_writes_reference_files(tmp_path):
    hf = tmp_path / "inv"
    hf.write_text("10.0.0.1\n10.0.0.2 extra\n# comment\n")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "cluster.py"), "get_hosts", "--hostfile", str(hf)],
                       cwd=tmp_path, capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    assert (tmp_path / "hosts").read_text().splitlines()[1] == "10.0.0.2 deeplearning-worker2"   # pytorch_ec2.py:656-708
    assert (tmp_path / "hosts_address").read_text().split() == ["10.0.0.1", "10.0.0.2"]
    assert (tmp_path / "hosts_alias").read_text().split()[0] == "deeplearning-worker1"


def test_cfg_interpolation():
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import cluster
    assert cluster.cfg["repo_dir"] == "/shared/erasurehead"


def test_makefile_targets_pass_13_positionals():
    p = subprocess.run(["make", "-n", "approx", "partialcyccoded"], cwd=ROOT, capture_output=True, text=True)
    assert p.returncode == 0
    for line in [l for l in p.stdout.splitlines() if "main.py" in l]:
        assert len(line.split("main.py")[1].split()) >= 13


def test_command_map_covers_the_reference(monkeypatch, tmp_path):
    """Every project-related entry of the reference's command map (pytorch_ec2.py:938-951) exists, and
    the composite clean_launch_and_run runs its steps in the reference's order."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import cluster
    ref = ["launch", "clean_launch_and_run", "shutdown", "get_hosts", "kill_all_python", "list_idle_instances",
           "list_running_instances", "kill_python", "run_command", "setup_nfs"]
    for name in ref:
        p = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "cluster.py"), name, "--help"],
                           capture_output=True, text=True)
        assert p.returncode == 0, (name, p.stderr)
    calls = []
    for fn in ("cmd_shutdown", "cmd_launch", "cmd_get_hosts", "cmd_bootstrap", "cmd_install", "cmd_torchrun"):
        monkeypatch.setattr(cluster, fn, lambda a, _n=fn: calls.append(_n))
    monkeypatch.setattr(cluster, "wait_until_running", lambda count, **kw: calls.append("wait"))
    monkeypatch.chdir(tmp_path)
    cluster.main(["clean_launch_and_run", "--count", "2", "--", "main.py", "9"])
    assert calls == ["cmd_shutdown", "cmd_launch", "wait", "cmd_get_hosts", "cmd_bootstrap", "cmd_install", "cmd_torchrun"]


class _FakeEc2:
    """Records the boto3 calls of the launch path; spot requests become active on the second poll."""

    def __init__(self):
        self.calls, self.polls = [], 0

    def request_spot_instances(self, **kw):
        self.calls.append(("request_spot_instances", kw))
        return {"SpotInstanceRequests": [{"SpotInstanceRequestId": f"sir-{i}"} for i in range(kw["InstanceCount"])]}

    def describe_spot_instance_requests(self, **kw):
        self.polls += 1
        ids = kw.get("SpotInstanceRequestIds") or ["sir-0"]
        done = self.polls >= 2
        return {"SpotInstanceRequests": [
            {"SpotInstanceRequestId": i, "State": "active" if done else "open", "InstanceId": "i-" + i[4:] if done else None,
             "Status": {"Code": "fulfilled" if done else "pending-evaluation", "Message": ""},
             "LaunchSpecification": {"KeyName": "k"}} for i in ids]}

    def create_tags(self, **kw):
        self.calls.append(("create_tags", kw))

    def run_instances(self, **kw):
        self.calls.append(("run_instances", kw))
        return {"Instances": [{"InstanceId": "i-x"}]}

    def cancel_spot_instance_requests(self, **kw):
        self.calls.append(("cancel_spot_instance_requests", kw))


def test_spot_launch_requests_waits_and_tags(monkeypatch):
    """The reference's default launch method (pytorch_ec2.py:29,39,176-205): request_spot_instances with the
    price as a string, poll until fulfilled, then tag; `ondemand` goes through run_instances."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import time
    import cluster
    fake = _FakeEc2()
    monkeypatch.setattr(cluster, "_ec2", lambda: fake)
    monkeypatch.setattr(time, "sleep", lambda s: None)
    monkeypatch.setitem(cluster.cfg, "key_name", "k")
    cluster.main(["launch", "--count", "3", "--method", "spot", "--spot-price", "12.5"])
    name, kw = fake.calls[0]
    assert name == "request_spot_instances" and kw["InstanceCount"] == 3 and kw["SpotPrice"] == "12.5"
    assert kw["LaunchSpecification"]["InstanceType"] == cluster.cfg["instance_type"]
    assert fake.polls == 2
    assert fake.calls[1][0] == "create_tags" and fake.calls[1][1]["Resources"] == ["i-0", "i-1", "i-2"]
    fake.calls.clear()
    cluster.main(["launch", "--count", "2", "--method", "ondemand"])
    assert fake.calls[0][0] == "run_instances" and fake.calls[0][1]["MinCount"] == 2
    fake.calls.clear()
    assert cluster.cancel_spot_requests(fake) == ["sir-0"]


def test_torchrun_starts_one_single_node_job_per_host():
    """Default: the multi-host launcher starts independent single-node jobs; {node} expands per host."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import cluster
    c0 = cluster.torchrun_command(0, 2, "/repo", 8, 29500, ["main.py", "9", "--seed={node}", "--output-dir=out{node}of{nodes}"])
    c1 = cluster.torchrun_command(1, 2, "/repo", 8, 29500, ["main.py", "9", "--seed={node}"])
    for c in (c0, c1):
        assert "--nnodes=1 " in c and "--master-addr 127.0.0.1" in c and "--node-rank" not in c
    assert "--seed=0" in c0 and "--output-dir=out0of2" in c0 and "--seed=1" in c1 and "run_node1.log" in c1


def test_torchrun_one_job_over_all_hosts():
    """--one-job: every host joins one rendezvous at the first host; EH_NODE_SIZE declares the node layout."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import cluster
    c0 = cluster.torchrun_command(0, 3, "/repo", 8, 29500, ["main.py", "25"], master="10.0.0.1")
    c2 = cluster.torchrun_command(2, 3, "/repo", 8, 29500, ["main.py", "25"], master="10.0.0.1")
    for i, c in ((0, c0), (2, c2)):
        assert f"--nnodes=3 --node-rank={i} " in c and "--master-addr 10.0.0.1" in c and "setsid env EH_NODE_SIZE=8 python" in c
    seen = []
    cluster_fan_out, cluster.fan_out = cluster.fan_out, lambda hosts, f: seen.extend(f(h) for h in hosts)
    try:
        hf = os.path.join(os.path.dirname(__file__), "_hosts_tmp")
        with open(hf, "w") as f:
            f.write("10.0.0.1\n10.0.0.2\n")
        cluster.main(["torchrun", "--hostfile", hf, "--gpus-per-node", "4", "--one-job", "--", "main.py", "9"])
    finally:
        cluster.fan_out = cluster_fan_out
        os.remove(hf)
    assert len(seen) == 2 and all("--nnodes=2" in c and "--master-addr 10.0.0.1" in c for c in seen)
    assert "--node-rank=1" in seen[1] and "EH_NODE_SIZE=4" in seen[1]


def test_grid_search_deals_points_to_hosts():
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import cluster
    pts = cluster.grid_points(["eta0=1,10", "num_collect=4,6,8"])
    assert len(pts) == 6 and pts[0] == {"eta0": "1", "num_collect": "4"} and pts[-1] == {"eta0": "10", "num_collect": "8"}
    rest = ["main.py", "9", "7680", "128", "/d/", "0", "x", "1", "1", "0", "3", "{num_collect}", "1", "AGD", "--eta0={eta0}"]
    cmds = cluster.grid_commands(4, "/repo", 8, 29500, rest, pts, "/shared/grid_search")
    assert len(cmds) == 4
    assert cmds[0].count("torch.distributed.run") == 2 and cmds[3].count("torch.distributed.run") == 1      # 6 points over 4 hosts
    assert "--eta0=1 " in cmds[0] and " 3 4 1 AGD" in cmds[0] and "/shared/grid_search/eta0_1_num_collect_4/" in cmds[0]
    assert "--output-dir=/shared/grid_search/eta0_10_num_collect_4/" in cmds[3]
    assert "{" not in "".join(cmds)
    seen = []
    keep, cluster.fan_out = cluster.fan_out, lambda hosts, f: seen.extend(f(h) for h in hosts)
    hf = os.path.join(os.path.dirname(__file__), "_hosts_tmp2")
    try:
        with open(hf, "w") as f:
            f.write("a\nb\n")
        cluster.main(["grid_search", "--hostfile", hf, "--grid", "seed=1,2,3", "--out-dir", "/o", "--", "main.py", "9", "--seed={seed}"])
    finally:
        cluster.fan_out = keep
        os.remove(hf)
    assert len(seen) == 2 and "--seed=1" in seen[0] and "--seed=3" in seen[0] and "--seed=2" in seen[1]


def test_check_env_script_runs_and_reports():
    """check_env (role of tools/pre_run.sh): the per-host script is valid shell + Python and names every item;
    on a machine without a GPU it must fail on the GPU line only where the stack is otherwise complete."""
    import subprocess
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import cluster
    p = subprocess.run(["bash", "-c", cluster.CHECK_ENV], capture_output=True, text=True, timeout=120)
    lines = p.stdout.strip().splitlines()
    names = [ln.split()[1].rstrip(":") for ln in lines]
    assert names[:1] == ["python"] and {"torch", "nvcc", "numpy", "scipy"} <= set(names), p.stdout + p.stderr
    assert all(ln.startswith(("ok   ", "FAIL ")) for ln in lines)
    assert (p.returncode == 0) == all(ln.startswith("ok") for ln in lines)


def test_dry_run_prints_commands_and_contacts_nobody(tmp_path, capsys):
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import cluster
    hf = tmp_path / "hosts"
    hf.write_text("hostA\nhostB\n")
    keep = cluster.ssh_run
    cluster.ssh_run = lambda *a, **k: (_ for _ in ()).throw(AssertionError("dry run must not open SSH connections"))
    try:
        cluster.main(["--dry-run", "torchrun", "--hostfile", str(hf), "--gpus-per-node", "4", "--one-job", "--", "main.py", "9"])
        cluster.main(["--dry-run", "check_env", str(hf)])
    finally:
        cluster.ssh_run = keep
        cluster.DRY_RUN = False
    out = capsys.readouterr().out
    assert out.count("would run:") == 4 and "[hostB] would run:" in out and "--node-rank=1" in out
    # the command of a host is valid shell (setsid must be followed by a program, not by an assignment)
    import subprocess
    line = [ln for ln in out.splitlines() if "torch.distributed.run" in ln][0].split("would run: ", 1)[1]
    assert subprocess.run(["bash", "-n", "-c", line]).returncode == 0 and "setsid env " in line
