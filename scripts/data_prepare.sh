#!/bin/bash
# Preprocess a real dataset into per-partition .npz files (role of the reference's data_prepare.sh).
N_PROCS=${N_PROCS:-31}
N_STRAGGLERS=${N_STRAGGLERS:-5}
N_PARTITIONS=${N_PARTITIONS:-30}
PARTIAL_CODED=${PARTIAL_CODED:-0}
DATA_FOLDER=${DATA_FOLDER:-./straggdata/}
DATASET=${DATASET:-kc_house_data}     # amazon-dataset | covtype | dna-dataset/dna | kc_house_data
cd "$(dirname "$0")/.."
exec python -m erasurehead_b200.data.arrange ${N_PROCS} ${DATA_FOLDER} ${DATASET} ${N_STRAGGLERS} ${N_PARTITIONS} ${PARTIAL_CODED}
