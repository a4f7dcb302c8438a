# Targets of the reference's Makefile / Makefile.make, now passing all 13 positionals (the
# reference's own targets pass 10 and only print the usage string).
N_PROCS ?= 9
N_GPUS ?= 1
N_STRAGGLERS ?= 1
N_PARTITIONS ?= 4
N_COLLECT ?= 6
ADD_DELAY ?= 1
UPDATE_RULE ?= AGD
PARTIAL_CODED ?= 0
DATA_FOLDER ?= ./straggdata/
IS_REAL ?= 0
DATASET ?= artificial
N_ROWS ?= 7680
N_COLS ?= 128
EXTRA ?=
PY ?= python
RUN = $(PY) -m torch.distributed.run --nnodes=1 --nproc-per-node $(N_GPUS) --master-addr 127.0.0.1 --master-port 29500 main.py
ARGS = $(N_PROCS) $(N_ROWS) $(N_COLS) $(DATA_FOLDER) $(IS_REAL) $(DATASET)
TAIL = $(N_COLLECT) $(ADD_DELAY) $(UPDATE_RULE) $(EXTRA)

build:
	bash csrc/build.sh
test:
	$(PY) -m pytest tests -q -m "not gpu"
test-gpu:
	$(PY) -m pytest tests -q -m gpu
test-parity-delays:      # the unmodified reference with its real Exp(0.5 s) sleeps against the model (CPU, ~2 min per scheme)
	EH_SLOW_TESTS=1 $(PY) -m pytest tests/test_reference_parity.py -q -s -k injected_delays
generate_random_data:
	$(PY) -m erasurehead_b200.data.generate $(N_PROCS) $(N_ROWS) $(N_COLS) $(DATA_FOLDER) $(N_STRAGGLERS) $(N_PARTITIONS) $(PARTIAL_CODED)
arrange_real_data:
	$(PY) -m erasurehead_b200.data.arrange $(N_PROCS) $(DATA_FOLDER) $(DATASET) $(N_STRAGGLERS) $(N_PARTITIONS) $(PARTIAL_CODED)
naive:
	$(RUN) $(ARGS) 0 $(N_STRAGGLERS) 0 0 $(TAIL)
cyccoded:
	$(RUN) $(ARGS) 1 $(N_STRAGGLERS) 0 0 $(TAIL)
repcoded:
	$(RUN) $(ARGS) 1 $(N_STRAGGLERS) 0 1 $(TAIL)
avoidstragg:
	$(RUN) $(ARGS) 1 $(N_STRAGGLERS) 0 2 $(TAIL)
approx:
	$(RUN) $(ARGS) 1 $(N_STRAGGLERS) 0 3 $(TAIL)
partialrepcoded:
	$(RUN) $(ARGS) 1 $(N_STRAGGLERS) $(N_PARTITIONS) 1 $(TAIL)
partialcyccoded:
	$(RUN) $(ARGS) 1 $(N_STRAGGLERS) $(N_PARTITIONS) 0 $(TAIL)
sanitize:
	bash tools/sanitize.sh
bench:
	$(PY) bench.py
.PHONY: build test test-gpu test-parity-delays generate_random_data arrange_real_data naive cyccoded repcoded avoidstragg approx partialrepcoded partialcyccoded sanitize bench
