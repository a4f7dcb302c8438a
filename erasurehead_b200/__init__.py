"""erasurehead_b200 -- Blackwell-native gradient coding (capabilities of ErasureHead).

Public API::

    from erasurehead_b200 import RunConfig, Engine, build_scheme
    cfg = RunConfig(n_procs=9, n_rows=..., n_cols=..., input_dir="", is_real=0, dataset="x",
                    is_coded=1, n_stragglers=1, partitions=0, coded_ver=3, num_collect=6,
                    add_delay=1, update_rule="AGD", synthetic=True)
    eng = Engine(cfg); res = eng.run(); ev = eng.evaluate(res)

``erasurehead_b200.compat`` has the reference's ten driver functions and its ``util.py`` helpers under their
own names and argument lists.
"""
from .config import RunConfig, parse_argv, scheme_from_triple, triple_from_scheme  # noqa: F401
from .schemes import SchemeSpec, build_scheme  # noqa: F401
from .engine import Engine  # noqa: F401
from . import codes, delays, oracle  # noqa: F401

__version__ = "0.1.0"
