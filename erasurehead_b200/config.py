"""Run configuration: the reference's 13-positional CLI plus ``--key=value`` overrides.

Positional order and meaning follow ``main.py:20-28`` of the reference exactly::

    n_procs n_rows n_cols input_dir is_real dataset is_coded n_stragglers
    partitions coded_ver num_collect add_delay update_rule

``n_procs`` keeps the reference meaning (workers + 1 master).  Here the master/decoder role
lives on GPU rank 0 next to its worker(s), so ``n_procs - 1`` workers are spread over the
available GPUs.  The scheme is selected by the triple ``(is_coded, partitions, coded_ver)``
as in ``main.py:62-92``; the literal dataset name ``kc_house_data`` switches to linear
regression for naive / replication / approx (``main.py:75,83,89``).

Hyper-parameters that the reference hard-codes in the source (``main.py:31-46``) are
overridable: ``--rounds``, ``--alpha``, ``--lr-schedule={const,t0decay,expdecay}``, ``--eta0``.
"""
from __future__ import annotations

import dataclasses
from dataclasses import dataclass
from typing import Optional, Sequence

import numpy as np

POSITIONALS = (
    "n_procs", "n_rows", "n_cols", "input_dir", "is_real", "dataset", "is_coded",
    "n_stragglers", "partitions", "coded_ver", "num_collect", "add_delay", "update_rule",
)

USAGE = ("Usage: python main.py " + " ".join(POSITIONALS) +
         " [--rounds=R --alpha=A --eta0=E --lr-schedule=const|t0decay|expdecay --seed=S "
         "--delay-scale=SEC --delay-mode=exp|fixed|dead|none --delay-carryover --gpus=N --synthetic "
         "--model=logistic|linear|softmax --classes=C --compat --output-dir=DIR "
         "--checkpoint-every=K --resume=PATH --eval=full|none --metrics-jsonl=PATH "
         "--synthetic-nnz=K --init=scheme|zeros|randn --dtype=fp32|bf16 --decoder=concurrent|inline --abandon "
         "--beta-push=p2p|multicast --decode=auto|table|solve --cuda-graph --debug-poison]")

SCHEMES = ("naive", "coded", "replication", "avoidstragg", "approx",
           "partial_coded", "partial_replication")


def scheme_from_triple(is_coded: int, partitions: int, coded_ver: int) -> str:
    """Dispatch tree of ``main.py:62-92``."""
    if not is_coded:
        return "naive"
    if partitions:
        if coded_ver == 1:
            return "partial_replication"
        if coded_ver == 0:
            return "partial_coded"
        raise ValueError(f"partial schemes need coded_ver in (0,1), got {coded_ver}")
    try:
        return {0: "coded", 1: "replication", 2: "avoidstragg", 3: "approx"}[coded_ver]
    except KeyError:
        raise ValueError(f"unknown coded_ver {coded_ver}") from None


def triple_from_scheme(scheme: str, partitions: int = 0) -> tuple:
    """Inverse of :func:`scheme_from_triple` (for launch scripts and tests)."""
    return {
        "naive": (0, 0, 1), "coded": (1, 0, 0), "replication": (1, 0, 1),
        "avoidstragg": (1, 0, 2), "approx": (1, 0, 3),
        "partial_coded": (1, partitions, 0), "partial_replication": (1, partitions, 1),
    }[scheme]


@dataclass
class RunConfig:
    # --- the 13 positionals ---
    n_procs: int
    n_rows: int
    n_cols: int
    input_dir: str
    is_real: int
    dataset: str
    is_coded: int
    n_stragglers: int
    partitions: int
    coded_ver: int
    num_collect: int
    add_delay: int
    update_rule: str
    # --- overrides (reference: edit-the-source constants, main.py:31-46) ---
    rounds: int = 100
    alpha: Optional[float] = None          # default 1/n_rows (main.py:34)
    eta0: float = 10.0                     # main.py:37
    lr_schedule: str = "const"             # const | t0decay (main.py:41-43) | expdecay (main.py:46)
    lr_t0: float = 90.0
    lr_decay: float = 0.98
    lr_values: Optional[Sequence[float]] = None   # explicit step per round (the reference's params[2]); overrides the schedule
    seed: Optional[int] = 0
    delay_scale: float = 0.5               # mean of the exponential delay in seconds (naive.py:147)
    delay_mode: str = "exp"                # exp | fixed | dead | none
    delay_fixed: float = 0.5               # sleep for the fixed-straggler-set variant (naive.py:143-145)
    delay_carryover: Optional[bool] = None  # a worker still inside its delay when a round ends starts the next round
                                           # late by what is left (what the reference's coded driver does: its master
                                           # never waits for stragglers, coded.py:137-166, so a sleeping worker stays
                                           # busy).  None = only with --compat for the coded scheme
    gpus: int = 0                          # 0 = all visible
    synthetic: bool = False                # generate data on the GPU instead of reading input_dir
    synthetic_nnz: int = 0                 # --synthetic with is_real=1: ones per row (0 = the named dataset's, data/shapes.py)
    model: Optional[str] = None            # None = reference rule (kc_house_data -> linear)
    classes: int = 1
    compat: bool = False                   # reproduce reference quirks Q1-Q3, Q9
    output_dir: Optional[str] = None
    checkpoint_every: int = 0
    resume: Optional[str] = None
    eval: str = "full"                     # full | none
    eval_precision: str = "tf32"           # tf32 (tcgen05 tensor cores) | fp32 (CUDA cores)
    metrics_jsonl: Optional[str] = None
    decode: str = "auto"                   # auto | table | solve
    quiet: bool = False
    cuda_graph: bool = False               # capture each enqueue of rounds into one CUDA graph
    beta_push: str = "p2p"                 # parameter broadcast: p2p (peer stores) | multicast (NVLS multimem.st)
    dtype: str = "fp32"                    # storage of dense features on the GPU: fp32 | bf16 (half the HBM traffic)
    decoder: str = "concurrent"            # concurrent: the decoder runs on its own high-priority stream on SMs the
                                           # gradient kernels leave free (a round costs max(compute, wait));
                                           # inline: after the root's gradient kernels on one stream (compute + wait)
    abandon: bool = False                  # dense single-column workers stop streaming X once the round they work for
                                           # has been decided without them (true skipping of a straggler's compute;
                                           # the reference can only cancel the send, coded.py:178-180, README.md:122)
    init: str = "scheme"                   # starting point: scheme (reference: randn for naive/replication/approx,
                                           # zeros otherwise -- SURVEY Q8) | zeros | randn
    debug_poison: bool = False             # fill every message slot with NaN before a run (stale-read detector)

    def __post_init__(self):
        if self.update_rule not in ("GD", "AGD"):
            raise ValueError("update_rule must be GD or AGD")  # naive.py:13
        if self.n_procs < 2:
            raise ValueError("n_procs must be >= 2 (one master + at least one worker)")
        if self.alpha is None:
            self.alpha = 1.0 / self.n_rows
        if self.input_dir and not self.input_dir.endswith("/"):
            self.input_dir += "/"
        if not self.is_real and not self.dataset.startswith("artificial-data/"):
            self.dataset = f"artificial-data/{self.n_rows}x{self.n_cols}"  # main.py:59-60
        if self.compat and self.scheme in ("avoidstragg", "partial_coded", "partial_replication"):
            # Q9: these drivers hard-wire AGD and have no delay block.
            self.update_rule = "AGD"
            self.add_delay = 0
        if self.resume is True or self.resume == "":
            raise ValueError("--resume needs the path of a checkpoint file (--resume=PATH)")
        if self.delay_carryover is None:
            self.delay_carryover = bool(self.compat and self.scheme == "coded")

    # ---- derived ----
    @property
    def n_workers(self) -> int:
        return self.n_procs - 1

    @property
    def scheme(self) -> str:
        return scheme_from_triple(self.is_coded, self.partitions, self.coded_ver)

    @property
    def model_name(self) -> str:
        if self.model:
            return self.model
        if self.dataset == "kc_house_data" and self.scheme in ("naive", "replication", "approx"):
            return "linear"
        return "logistic"

    @property
    def n_partitions(self) -> int:
        """Number of partition files on disk."""
        if self.scheme.startswith("partial_"):
            return (self.partitions - self.n_stragglers) * self.n_workers  # main.py:66
        return self.n_workers

    @property
    def data_dir(self) -> str:
        """Directory layout of ``main.py:66-72``."""
        base = self.input_dir + self.dataset
        if self.scheme.startswith("partial_"):
            return f"{base}/partial/{self.n_partitions}/"
        return f"{base}/{self.n_workers}/"

    @property
    def results_dir(self) -> str:
        return self.output_dir or (self.data_dir + "results/")

    def learning_rates(self) -> np.ndarray:
        """The three schedules present (two of them commented out) in ``main.py:37-46``."""
        R = self.rounds
        if self.lr_values is not None:
            v = np.asarray(self.lr_values, dtype=np.float64).reshape(-1)
            if v.size < R:
                raise ValueError(f"lr_values has {v.size} entries for {R} rounds")
            return v[:R].copy()
        if self.lr_schedule == "const":
            return self.eta0 * np.ones(R)
        if self.lr_schedule == "t0decay":
            return np.array([self.eta0 * self.lr_t0 / (i + self.lr_t0) for i in range(1, R + 1)])
        if self.lr_schedule == "expdecay":
            return np.array([self.eta0 * self.lr_decay ** i for i in range(1, R + 1)])
        raise ValueError(f"unknown lr schedule {self.lr_schedule}")

    def replace(self, **kw) -> "RunConfig":
        return dataclasses.replace(self, **kw)


_INT_FIELDS = {"n_procs", "n_rows", "n_cols", "is_real", "is_coded", "n_stragglers",
               "partitions", "coded_ver", "num_collect", "add_delay"}


def parse_argv(argv: Sequence[str]) -> RunConfig:
    """Parse ``argv`` (without the program name).  The first 13 tokens are the reference
    positionals; anything after must be ``--key=value`` / ``--flag``."""
    pos = [a for a in argv if not a.startswith("--")]
    opts = [a for a in argv if a.startswith("--")]
    if len(pos) != len(POSITIONALS):
        raise SystemExit(USAGE)
    kw = {}
    for name, val in zip(POSITIONALS, pos):
        kw[name] = int(val) if name in _INT_FIELDS else val
    fields = {f.name: f for f in dataclasses.fields(RunConfig)}
    for o in opts:
        key, _, val = o[2:].partition("=")
        key = key.replace("-", "_")
        if key not in fields or key in POSITIONALS:
            raise SystemExit(f"unknown option --{key}\n{USAGE}")
        typ = fields[key].type
        if val == "":
            kw[key] = True
        elif "int" in str(typ):
            kw[key] = int(val)
        elif "float" in str(typ):
            kw[key] = float(val)
        elif "bool" in str(typ):
            kw[key] = val.lower() in ("1", "true", "yes", "on")
        else:
            kw[key] = val
    return RunConfig(**kw)
