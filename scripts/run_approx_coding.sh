#!/bin/bash
# Launch one training run (role of the reference's run_approx_coding.sh: a variable block and
# one launcher line).  torchrun starts one process per GPU; N_PROCS keeps the reference meaning
# (workers + 1), the workers are spread over the GPUs.
N_PROCS=${N_PROCS:-31}            # 30 logical workers (run_approx_coding.sh:2)
N_GPUS=${N_GPUS:-8}
N_STRAGGLERS=${N_STRAGGLERS:-3}
N_COLLECT=${N_COLLECT:-15}
UPDATE_RULE=${UPDATE_RULE:-AGD}
N_PARTITIONS=${N_PARTITIONS:-30}  # partial schemes only
ADD_DELAY=${ADD_DELAY:-0}
DATA_FOLDER=${DATA_FOLDER:-./straggdata/}
IS_REAL=${IS_REAL:-1}
#DATASET=covtype        N_ROWS=396112 N_COLS=15509
#DATASET=amazon-dataset N_ROWS=26215  N_COLS=241915
DATASET=${DATASET:-kc_house_data}
N_ROWS=${N_ROWS:-17290}
N_COLS=${N_COLS:-27654}
# MODES (is_coded partitions coded_ver):  1 0 1 grad coding rep | 1 0 3 approx coding | 0 x x vanilla GD
# | 1 0 0 cyclic MDS | 1 0 2 ignore stragglers | 1 P 0 partial coded | 1 P 1 partial replication
MODE=${MODE:-"1 0 3"}
read IS_CODED PARTS CODED_VER <<< "$MODE"
HOSTFILE=${HOSTFILE:-}            # optional: one host per line (the reference's hosts_address, run_approx_coding.sh:47-49)
ONE_JOB=${ONE_JOB:-1}             # with HOSTFILE: 1 = one job over all hosts, as mpirun --hostfile did (first host's GPUs
                                  # exchange through peer memory, the others over NCCL); 0 = one independent job per host
EXTRA="$@"

cd "$(dirname "$0")/.."
if [ -n "$HOSTFILE" ]; then
  [ "$ONE_JOB" = "1" ] && ONE="--one-job" || ONE=""
  exec python tools/cluster.py torchrun --hostfile "$HOSTFILE" --gpus-per-node ${N_GPUS} ${ONE} -- \
    main.py ${N_PROCS} ${N_ROWS} ${N_COLS} ${DATA_FOLDER} ${IS_REAL} ${DATASET} ${IS_CODED} ${N_STRAGGLERS} ${PARTS} ${CODED_VER} ${N_COLLECT} ${ADD_DELAY} ${UPDATE_RULE} ${EXTRA}
fi
exec python -m torch.distributed.run --nnodes=1 --nproc-per-node ${N_GPUS} --master-addr 127.0.0.1 --master-port ${MASTER_PORT:-29500} \
  main.py ${N_PROCS} ${N_ROWS} ${N_COLS} ${DATA_FOLDER} ${IS_REAL} ${DATASET} ${IS_CODED} ${N_STRAGGLERS} ${PARTS} ${CODED_VER} ${N_COLLECT} ${ADD_DELAY} ${UPDATE_RULE} ${EXTRA}
