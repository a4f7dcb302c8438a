#!/usr/bin/env python
"""Single-node launcher: ``tools/launch.py --gpus N -- <13 positionals> [--overrides]``
(torchrun with the loopback rendezvous this environment needs)."""
import subprocess
import sys

if __name__ == "__main__":
    argv = sys.argv[1:]
    gpus = 1
    if argv[:1] == ["--gpus"]:
        gpus, argv = int(argv[1]), argv[2:]
    if argv[:1] == ["--"]:
        argv = argv[1:]
    sys.exit(subprocess.call([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                              f"--nproc-per-node={gpus}", "--master-addr", "127.0.0.1", "--master-port", "29500",
                              "main.py"] + argv))
