#!/usr/bin/env python
"""Cluster operations tool (role of the reference's ``tools/pytorch_ec2.py`` + shell helpers).

Sub-commands mirror the reference's command map (pytorch_ec2.py:938-951) where they still make
sense for GPU nodes:

  launch [--method spot|ondemand] / shutdown / list_running_instances
                                               EC2 instances through boto3 (only if boto3 is installed).  Spot
                                               (the reference's default, pytorch_ec2.py:29,39,176-205): one request of
                                               `--count` instances at `--spot-price`, wait until every request is
                                               fulfilled, tag the instances; shutdown also cancels open spot requests
                                               (terminate_all_requests, :129-154)
  get_hosts                                    write ``hosts``, ``hosts_alias``, ``hosts_address`` (:656-708)
  run_command  <hostfile> -- <cmd>             run a shell command on every host (paramiko SSH fan-out, :259-299)
  list_idle_instances <hostfile>               hosts with no GPU compute process (:304-336)
  kill_python <hostfile>                       stop the training processes this tool started (by recorded PID)
  setup_nfs <hostfile> <export_dir>            export a directory from the first host and mount it on the rest (:880-900)
  install <hostfile>                           build the CUDA library on every host (tools/install.sh role)
  bootstrap <hostfile>                         node set-up of tools/local_script.sh / remote_script.sh: append the
                                               `hosts` aliases to /etc/hosts, share one cluster SSH key pair, accept
                                               host keys, so that every node reaches every other by alias
  update_repo <hostfile>                       `git pull` in the repository on every host (tools/update_git_dir.sh role)
  check_env <hostfile>                         role of tools/pre_run.sh, checked instead of installed: Python, torch + CUDA,
                                               sm_100 GPUs, NCCL, nvcc, numpy / scipy on every host
  clean_launch_and_run --count N -- main.py <13 positionals>
                                               shutdown, launch, wait until every instance is up and reachable,
                                               get_hosts, bootstrap, install, torchrun (pytorch_ec2.py:938 map entry)
  torchrun --hostfile F --gpus-per-node G [--one-job] -- main.py <13 positionals>
                                               default: one INDEPENDENT single-node ``torch.distributed.run`` job per
                                               host; ``{node}`` / ``{nodes}`` in the arguments expand to the host's
                                               index / the host count (seeds, output directories).
                                               --one-job: ONE job over all hosts, as ``mpirun --hostfile`` ran the
                                               reference: the first host's GPUs form the peer-memory domain, the
                                               other hosts' GPUs are reached over NCCL (parallel/bridge.py)

  grid_search --hostfile F --grid name=v1,v2 [--grid ...] --out-dir D -- main.py <13 positionals> --opt={name}
                                               hyper-parameter grid dealt round-robin to the hosts, each host runs its
                                               points one after the other and writes results + log under D/<point>/
                                               (role of run_mxnet_grid_search / run_mxnet_loss_curve, :374-652: the
                                               per-iteration loss curves are the result files of every run)

``--dry-run`` before the sub-command prints the per-host commands instead of running them.

Unlike the reference, processes are never killed by pattern (``pkill python``): the launcher
records the PID it started on every host under ``~/.erasurehead_b200/pids`` and only those are
signalled.  Credentials come from the SSH agent / ``~/.ssh/config``; nothing is hard-coded.
"""
from __future__ import annotations

import argparse
import os
import shlex
import sys
from concurrent.futures import ThreadPoolExecutor
from typing import List

PID_DIR = "~/.erasurehead_b200/pids"


class Cfg(dict):
    """Dict with ``%(key)s`` self-interpolation, like the reference's ``Cfg`` (pytorch_ec2.py:12-20)."""

    def __getitem__(self, k):
        v = dict.__getitem__(self, k)
        return v % self if isinstance(v, str) and "%(" in v else v


cfg = Cfg({
    "name": "erasurehead", "region": "us-west-2", "availability_zone": "us-west-2b",
    "instance_type": "p6-b200.48xlarge", "image_id": "", "key_name": "", "ssh_username": "ubuntu",
    "n_workers": 1, "nfs_mount_point": "/shared", "repo_dir": "%(nfs_mount_point)s/%(name)s",
    "method": "ondemand",          # "spot" is the reference's default (pytorch_ec2.py:29); GPU capacity is usually reserved
    "spot_price": "",              # max price per instance-hour as a string (pytorch_ec2.py:39); "" = on-demand price cap
    "security_groups": ["default"],
})


def read_hosts(path: str) -> List[str]:
    with open(path) as f:
        return [ln.split()[0] for ln in f if ln.strip() and not ln.startswith("#")]


def ssh_run(host: str, command: str, user: str = None, timeout: int = 600):
    import paramiko
    cli = paramiko.SSHClient()
    cli.set_missing_host_key_policy(paramiko.AutoAddPolicy())
    cli.connect(host, username=user or os.environ.get("EH_SSH_USER"), timeout=30)
    try:
        _, out, err = cli.exec_command(command, timeout=timeout)
        rc = out.channel.recv_exit_status()
        return rc, out.read().decode(), err.read().decode()
    finally:
        cli.close()


DRY_RUN = False          # --dry-run: print what would be run on which host, contact nobody


def fan_out(hosts: List[str], command, user=None):
    if DRY_RUN:
        res = {}
        for h in hosts:
            print(f"[{h}] would run: {command(h) if callable(command) else command}")
            res[h] = (0, "", "")
        return res
    with ThreadPoolExecutor(max_workers=min(32, max(1, len(hosts)))) as ex:
        futs = {h: ex.submit(ssh_run, h, command(h) if callable(command) else command, user) for h in hosts}
        res = {}
        for h, f in futs.items():
            try:
                res[h] = f.result()
            except Exception as e:  # noqa: BLE001
                res[h] = (255, "", str(e))
    for h, (rc, out, err) in res.items():
        print(f"[{h}] rc={rc}\n{out}{err}".rstrip())
    return res


def _ec2():
    try:
        import boto3
    except ImportError:
        sys.exit("boto3 is not installed: EC2 sub-commands are unavailable here (use a static hostfile)")
    return boto3.client("ec2", region_name=cfg["region"])


def launch_spec(count: int) -> dict:
    """Arguments shared by both launch methods (the reference's `launch_specs`, pytorch_ec2.py:184-188)."""
    return {"ImageId": cfg["image_id"], "InstanceType": cfg["instance_type"], "KeyName": cfg["key_name"],
            "Placement": {"AvailabilityZone": cfg["availability_zone"]}, "SecurityGroups": list(cfg["security_groups"])}


def spot_request_args(count: int, price: str) -> dict:
    args = {"InstanceCount": count, "Type": "one-time", "LaunchSpecification": launch_spec(count)}
    if price:
        args["SpotPrice"] = str(price)
    return args


def wait_spot_fulfilled(ec2, request_ids, timeout_s: int = 1800, poll_s: int = 10, sleep=None):
    """Poll the spot requests until every one is `active` (fulfilled); print state - status lines like the
    reference's wait_until_instance_request_status_fulfilled (pytorch_ec2.py:232-256).  A request that
    ends `closed` / `cancelled` / `failed` aborts.  Returns the instance ids."""
    import time
    sleep = sleep or time.sleep
    t0 = time.time()
    while True:
        r = ec2.describe_spot_instance_requests(SpotInstanceRequestIds=list(request_ids))
        reqs = r["SpotInstanceRequests"]
        for q in reqs:
            st = q.get("Status", {})
            print(f"{q['SpotInstanceRequestId']} {q['State']} - {st.get('Code', '')}: {st.get('Message', '')}")
        bad = [q for q in reqs if q["State"] in ("closed", "cancelled", "failed")]
        if bad:
            sys.exit(f"spot request {bad[0]['SpotInstanceRequestId']} ended {bad[0]['State']}: "
                     f"{bad[0].get('Status', {}).get('Message', '')}")
        if reqs and all(q["State"] == "active" and q.get("InstanceId") for q in reqs):
            return [q["InstanceId"] for q in reqs]
        if time.time() - t0 > timeout_s:
            sys.exit(f"spot requests not fulfilled within {timeout_s} s")
        sleep(poll_s)


def cancel_spot_requests(ec2):
    """Cancel this cluster's open/active spot requests (terminate_all_requests, pytorch_ec2.py:129-154);
    requests are recognised by the launch specification's key name, as in the reference."""
    r = ec2.describe_spot_instance_requests(Filters=[{"Name": "state", "Values": ["open", "active"]}])
    ids = [q["SpotInstanceRequestId"] for q in r["SpotInstanceRequests"]
           if q.get("LaunchSpecification", {}).get("KeyName") == cfg["key_name"]]
    if ids:
        ec2.cancel_spot_instance_requests(SpotInstanceRequestIds=ids)
    print("cancelled spot requests", ids)
    return ids


def cmd_launch(a):
    ec2 = _ec2()
    method = getattr(a, "method", None) or cfg["method"]
    tags = [{"Key": "Name", "Value": cfg["name"]}]
    if method == "spot":
        price = getattr(a, "spot_price", None) or cfg["spot_price"]
        r = ec2.request_spot_instances(**spot_request_args(a.count, price))
        req_ids = [q["SpotInstanceRequestId"] for q in r["SpotInstanceRequests"]]
        print("spot requests", req_ids)
        ids = wait_spot_fulfilled(ec2, req_ids)
        ec2.create_tags(Resources=ids, Tags=tags)        # spot launch specifications cannot carry instance tags
        print("launched (spot)", ids)
    elif method in ("ondemand", "reserved"):
        r = ec2.run_instances(MinCount=a.count, MaxCount=a.count,
                              TagSpecifications=[{"ResourceType": "instance", "Tags": tags}], **launch_spec(a.count))
        print("launched", [i["InstanceId"] for i in r["Instances"]])
    else:
        sys.exit(f"unknown launch method {method!r} (spot | ondemand)")


def wait_until_running(count: int, timeout_s: int = 900, poll_s: int = 10):
    """Block until `count` tagged instances are running and answer on port 22 (the reference's wait
    loops, pytorch_ec2.py:209-256)."""
    import socket
    import time
    t0 = time.time()
    while time.time() - t0 < timeout_s:
        inst = _instances()
        if len(inst) >= count:
            up = 0
            for i in inst:
                try:
                    socket.create_connection((i.get("PublicIpAddress") or i["PrivateIpAddress"], 22), timeout=3).close()
                    up += 1
                except OSError:
                    pass
            if up >= count:
                return inst
        time.sleep(poll_s)
    sys.exit(f"only {len(_instances())} of {count} instances came up within {timeout_s} s")


def _instances(states=("running",)):
    ec2 = _ec2()
    r = ec2.describe_instances(Filters=[{"Name": "tag:Name", "Values": [cfg["name"]]},
                                        {"Name": "instance-state-name", "Values": list(states)}])
    return [i for res in r["Reservations"] for i in res["Instances"]]


def cmd_shutdown(a):
    cancel_spot_requests(_ec2())
    ids = [i["InstanceId"] for i in _instances(("running", "pending"))]
    if ids:
        _ec2().terminate_instances(InstanceIds=ids)
    print("terminated", ids)


def cmd_list_running(a):
    for i in _instances():
        print(i["InstanceId"], i.get("PublicIpAddress"), i.get("PrivateIpAddress"))


def cmd_get_hosts(a):
    """hosts: 'ip alias' lines for /etc/hosts; hosts_alias: aliases; hosts_address: private ips."""
    if a.hostfile:
        ips = read_hosts(a.hostfile)
    else:
        ips = [i["PrivateIpAddress"] for i in _instances()]
    with open("hosts", "w") as h, open("hosts_alias", "w") as al, open("hosts_address", "w") as ad:
        for k, ip in enumerate(ips):
            alias = f"deeplearning-worker{k + 1}"
            h.write(f"{ip} {alias}\n"); al.write(alias + "\n"); ad.write(ip + "\n")
    print(f"wrote hosts, hosts_alias, hosts_address ({len(ips)} nodes)")


def cmd_run_command(a):
    fan_out(read_hosts(a.hostfile), " ".join(shlex.quote(x) for x in a.rest))


def cmd_list_idle(a):
    res = fan_out(read_hosts(a.hostfile),
                  "nvidia-smi --query-compute-apps=pid --format=csv,noheader | wc -l")
    print("idle:", [h for h, (rc, out, _) in res.items() if rc == 0 and out.strip() == "0"])


def cmd_kill_python(a):
    fan_out(read_hosts(a.hostfile),
            f"for f in {PID_DIR}/*.pid; do [ -f \"$f\" ] && kill -- -$(cat \"$f\") 2>/dev/null; rm -f \"$f\"; done; true")


def cmd_setup_nfs(a):
    hosts = read_hosts(a.hostfile)
    head, rest = hosts[0], hosts[1:]
    d = a.export_dir
    fan_out([head], f"sudo mkdir -p {d} && echo '{d} *(rw,sync,no_subtree_check,no_root_squash)' | "
                    "sudo tee -a /etc/exports && sudo exportfs -ra")
    fan_out(rest, f"sudo mkdir -p {d} && sudo mount -t nfs {head}:{d} {d}")


def cmd_install(a):
    repo = a.repo or cfg["repo_dir"]
    fan_out(read_hosts(a.hostfile), f"cd {repo} && bash csrc/build.sh | tail -1")


CHECK_ENV = ("python - <<'PY'\n"
             "import sys, shutil, subprocess\n"
             "ok = True\n"
             "def say(k, v, good=True):\n"
             "    global ok; ok &= bool(good); print(('ok   ' if good else 'FAIL ') + k + ': ' + str(v))\n"
             "say('python', sys.version.split()[0], sys.version_info >= (3, 10))\n"
             "try:\n"
             "    import torch\n"
             "    say('torch', torch.__version__ + ' cuda ' + str(torch.version.cuda), torch.cuda.is_available())\n"
             "    n = torch.cuda.device_count()\n"
             "    caps = sorted({torch.cuda.get_device_capability(i) for i in range(n)})\n"
             "    say('gpus', str(n) + ' x ' + (torch.cuda.get_device_name(0) if n else '-') + ' ' + str(caps), n > 0 and caps == [(10, 0)])\n"
             "    import torch.distributed as d\n"
             "    say('nccl', '.'.join(map(str, torch.cuda.nccl.version())), d.is_nccl_available())\n"
             "except Exception as e:\n"
             "    say('torch', repr(e), False)\n"
             "nv = shutil.which('nvcc') or '/usr/local/cuda/bin/nvcc'\n"
             "try:\n"
             "    out = [l for l in subprocess.run([nv, '--version'], capture_output=True, text=True).stdout.splitlines() if 'release' in l]\n"
             "    say('nvcc', out[0].strip() if out else 'no version line', bool(out) and ('release 12.' in out[0] or 'release 13.' in out[0]))\n"
             "except Exception as e:\n"
             "    say('nvcc', repr(e), False)\n"
             "for m in ('numpy', 'scipy'):\n"
             "    try:\n"
             "        say(m, __import__(m).__version__)\n"
             "    except Exception as e:\n"
             "        say(m, repr(e), False)\n"
             "sys.exit(0 if ok else 1)\n"
             "PY")


def cmd_check_env(a):
    """What tools/pre_run.sh was for -- a node that can run the code -- checked instead of installed (the
    software stack is the machine image): Python, torch with CUDA, sm_100 GPUs, NCCL, nvcc, numpy / scipy on every
    host; a host that fails returns rc 1."""
    res = fan_out(read_hosts(a.hostfile), CHECK_ENV)
    bad = [h for h, (rc, _, _) in res.items() if rc != 0]
    if bad:
        sys.exit(f"environment check failed on: {' '.join(bad)}")


def cmd_bootstrap(a):
    """/etc/hosts aliases + one shared key pair + known_hosts on every node (tools/local_script.sh,
    tools/remote_script.sh).  Reads the `hosts` file written by get_hosts from the current directory."""
    hosts = read_hosts(a.hostfile)
    lines = open("hosts").read().strip().replace("'", "") if os.path.exists("hosts") else ""
    head = hosts[0]
    # the head node creates the cluster key once; everyone authorises it and learns the aliases
    fan_out([head], "test -f ~/.ssh/id_cluster || ssh-keygen -q -t ed25519 -N '' -f ~/.ssh/id_cluster; cat ~/.ssh/id_cluster.pub")
    rc, pub, _ = ssh_run(head, "cat ~/.ssh/id_cluster.pub")
    pub = pub.strip()
    fan_out(hosts, f"mkdir -p ~/.ssh && chmod 700 ~/.ssh && grep -qF '{pub}' ~/.ssh/authorized_keys 2>/dev/null || "
                   f"echo '{pub}' >> ~/.ssh/authorized_keys; "
                   + (f"printf '%s\n' '{lines}' | while read ip al; do grep -q \"$al\" /etc/hosts || "
                      f"echo \"$ip $al\" | sudo tee -a /etc/hosts >/dev/null; done; " if lines else "")
                   + "for h in " + " ".join(hosts) + "; do ssh-keyscan -H $h >> ~/.ssh/known_hosts 2>/dev/null; done; true")


def cmd_update_repo(a):
    repo = a.repo or cfg["repo_dir"]
    fan_out(read_hosts(a.hostfile), f"cd {repo} && git pull --ff-only | tail -1")


def cmd_clean_launch_and_run(a):
    """shutdown -> launch -> wait -> get_hosts -> bootstrap -> install -> torchrun."""
    cmd_shutdown(a)
    cmd_launch(a)
    wait_until_running(a.count)
    a.hostfile = None
    cmd_get_hosts(a)
    a.hostfile = "hosts_address"
    cmd_bootstrap(a)
    cmd_install(a)
    cmd_torchrun(a)


def torchrun_command(host_index: int, n_hosts: int, repo: str, gpus_per_node: int, port: int, rest: List[str],
                     master: str = "") -> str:
    """Shell command for ONE host.

    Default (``master`` empty): an independent single-node job per host.  ``{node}`` and ``{nodes}`` in the
    arguments become the host's index and the host count, so that hosts can run different seeds /
    configurations / output directories.
    With ``master`` (the first host's address): this host's share of ONE job over all hosts -- what
    ``mpirun --hostfile`` did for the reference (run_approx_coding.sh:47-49).  Rank 0's node is the
    peer-memory domain, the GPUs of the other hosts are reached through the NCCL bridge
    (erasurehead_b200/parallel/bridge.py); EH_NODE_SIZE tells every rank the layout without a collective."""
    args = " ".join(shlex.quote(x.replace("{node}", str(host_index)).replace("{nodes}", str(n_hosts))) for x in rest)
    if master:
        rdzv = (f"env EH_NODE_SIZE={gpus_per_node} python -m torch.distributed.run --nnodes={n_hosts} --node-rank={host_index} "
                f"--nproc-per-node={gpus_per_node} --master-addr {master} --master-port {port}")
    else:
        rdzv = (f"python -m torch.distributed.run --nnodes=1 --nproc-per-node={gpus_per_node} "
                f"--master-addr 127.0.0.1 --master-port {port}")
    return (f"mkdir -p {PID_DIR} && cd {repo} && setsid {rdzv} {args} "
            f"> run_node{host_index}.log 2>&1 & echo $! > {PID_DIR}/run.pid; wait; tail -5 run_node{host_index}.log")


def cmd_torchrun(a):
    hosts = read_hosts(a.hostfile)
    repo = a.repo or os.getcwd()
    master = hosts[0] if getattr(a, "one_job", False) and len(hosts) > 1 else ""
    fan_out(hosts, lambda h: torchrun_command(hosts.index(h), len(hosts), repo, a.gpus_per_node, a.port, a.rest, master))


def grid_points(specs: List[str]) -> List[dict]:
    """``["eta0=1,10", "n_collect=4,6"]`` -> the Cartesian product as a list of {name: value} dicts."""
    import itertools
    names, values = [], []
    for sp in specs:
        name, _, vals = sp.partition("=")
        if not name or not vals:
            raise SystemExit(f"--grid wants name=v1,v2,... (got {sp!r})")
        names.append(name)
        values.append(vals.split(","))
    return [dict(zip(names, combo)) for combo in itertools.product(*values)]


def grid_commands(n_hosts: int, repo: str, gpus_per_node: int, port: int, rest: List[str], points: List[dict],
                  out_dir: str) -> List[str]:
    """One shell command per host: the grid points are dealt round-robin to the hosts and every host runs its
    share one after the other as single-node jobs (the role of the reference tool's ``run_mxnet_grid_search`` /
    ``run_mxnet_loss_curve``, pytorch_ec2.py:374-652: host k works on slice k of the grid and writes under the
    shared directory).  ``{name}`` in the arguments is replaced by the point's value; the result files of a point
    go to ``<out_dir>/<name>_<value>_...`` (``--output-dir``), its log next to them."""
    cmds = []
    for h in range(n_hosts):
        runs = []
        for pt in points[h::n_hosts]:
            tag = "_".join(f"{k}_{v}" for k, v in pt.items())
            args = [x for x in rest]
            for k, v in pt.items():
                args = [a.replace("{" + k + "}", v) for a in args]
            d = f"{out_dir.rstrip('/')}/{tag}"
            argstr = " ".join(shlex.quote(a) for a in args + [f"--output-dir={d}/"])
            runs.append(f"mkdir -p {shlex.quote(d)} && python -m torch.distributed.run --nnodes=1 "
                        f"--nproc-per-node={gpus_per_node} --master-addr 127.0.0.1 --master-port {port} {argstr} "
                        f"> {shlex.quote(d)}/run.log 2>&1")
        body = " ; ".join(runs) if runs else "true"
        cmds.append(f"mkdir -p {PID_DIR} && cd {repo} && (setsid bash -c {shlex.quote(body)} & echo $! > {PID_DIR}/run.pid; wait)")
    return cmds


def cmd_grid_search(a):
    hosts = read_hosts(a.hostfile)
    repo = a.repo or os.getcwd()
    points = grid_points(a.grid)
    cmds = grid_commands(len(hosts), repo, a.gpus_per_node, a.port, a.rest, points, a.out_dir or cfg["nfs_mount_point"] + "/grid_search")
    print(f"{len(points)} grid points over {len(hosts)} hosts")
    fan_out(hosts, lambda h: cmds[hosts.index(h)])


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--dry-run", action="store_true", help="print the per-host commands instead of running them over SSH")
    sub = ap.add_subparsers(dest="cmd", required=True)
    p = sub.add_parser("launch"); p.add_argument("--count", type=int, default=cfg["n_workers"])
    p.add_argument("--method", choices=["spot", "ondemand", "reserved"]); p.add_argument("--spot-price")
    p.set_defaults(fn=cmd_launch)
    sub.add_parser("shutdown").set_defaults(fn=cmd_shutdown)
    sub.add_parser("list_running_instances").set_defaults(fn=cmd_list_running)
    p = sub.add_parser("get_hosts"); p.add_argument("--hostfile"); p.set_defaults(fn=cmd_get_hosts)
    for name, fn in (("run_command", cmd_run_command), ("list_idle_instances", cmd_list_idle),
                     ("kill_python", cmd_kill_python), ("kill_all_python", cmd_kill_python), ("install", cmd_install),
                     ("bootstrap", cmd_bootstrap), ("update_repo", cmd_update_repo), ("check_env", cmd_check_env)):
        p = sub.add_parser(name); p.add_argument("hostfile"); p.add_argument("--repo")
        p.add_argument("rest", nargs=argparse.REMAINDER); p.set_defaults(fn=fn)
    p = sub.add_parser("setup_nfs"); p.add_argument("hostfile"); p.add_argument("export_dir"); p.set_defaults(fn=cmd_setup_nfs)
    p = sub.add_parser("torchrun"); p.add_argument("--hostfile", required=True)
    p.add_argument("--gpus-per-node", type=int, default=8); p.add_argument("--port", type=int, default=29500)
    p.add_argument("--one-job", action="store_true", help="one job over all hosts instead of one job per host")
    p.add_argument("--repo"); p.add_argument("rest", nargs=argparse.REMAINDER); p.set_defaults(fn=cmd_torchrun)
    p = sub.add_parser("grid_search"); p.add_argument("--hostfile", required=True)
    p.add_argument("--grid", action="append", default=[], help="name=v1,v2,... (repeatable); {name} in the arguments")
    p.add_argument("--out-dir", default=""); p.add_argument("--gpus-per-node", type=int, default=8)
    p.add_argument("--port", type=int, default=29500); p.add_argument("--repo")
    p.add_argument("rest", nargs=argparse.REMAINDER); p.set_defaults(fn=cmd_grid_search)
    p = sub.add_parser("clean_launch_and_run"); p.add_argument("--count", type=int, default=cfg["n_workers"])
    p.add_argument("--method", choices=["spot", "ondemand", "reserved"]); p.add_argument("--spot-price")
    p.add_argument("--gpus-per-node", type=int, default=8); p.add_argument("--port", type=int, default=29500)
    p.add_argument("--repo"); p.add_argument("rest", nargs=argparse.REMAINDER); p.set_defaults(fn=cmd_clean_launch_and_run)
    a = ap.parse_args(argv)
    global DRY_RUN
    DRY_RUN = bool(a.dry_run)
    if getattr(a, "rest", None) and a.rest and a.rest[0] == "--":
        a.rest = a.rest[1:]
    a.fn(a)


if __name__ == "__main__":
    main()
