# The reference ships two variable blocks for the same targets: Makefile (11 processes, s = 4, P = 10) and
# Makefile.make (21 processes, s = 5, P = 20; Makefile.make:2-8).  This is the second one: the same targets as
# ./Makefile with that cluster shape as defaults (20 workers spread over the visible GPUs).
#     make -f Makefile.make approx N_GPUS=8
N_PROCS ?= 21
N_STRAGGLERS ?= 4
N_PARTITIONS ?= 20
N_COLLECT ?= 10
N_ROWS ?= 8000
include Makefile
