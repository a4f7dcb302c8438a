#!/bin/bash
# sweep of the one-pass multi-column kernel variants (cluster size) against the two-pass kernels
S="2097152,512,16 1048576,1024,16 524288,2048,16"
f() { python benchmarks/kernel_bench.py --only mc --iters 8 --mc-shapes $S 2>&1 | grep kernel | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print('$1', r['d'], 'ms', round(r['ms'],3), 'frac', round([v for k,v in r.items() if k.startswith('frac')][0],3))"; }
EH_MC_TWOPASS=1 f twopass
f auto
EH_MC_ONEPASS=cluster EH_MC_CS=2 f cs2
EH_MC_ONEPASS=cluster EH_MC_CS=4 f cs4
EH_MC_ONEPASS=cluster EH_MC_CS=8 f cs8
