#!/usr/bin/env python
"""Entry point with the reference's CLI (main.py of Distributed-Deep-Learning/ErasureHead).

    torchrun --nproc-per-node 8 main.py 9 <rows> <cols> <dir> 0 x 1 1 0 3 6 1 AGD --synthetic
"""
import sys

from erasurehead_b200.cli import main

if __name__ == "__main__":
    sys.exit(main())
