import sys; sys.path.insert(0, '.')
import torch
from erasurehead_b200 import Engine, RunConfig, triple_from_scheme
ic, pp, cv = triple_from_scheme("approx")
cfg = RunConfig(9, 8 * 256, 4096, "", 0, "x", ic, 1, pp, cv, 4, 0, "AGD", rounds=40, synthetic=True, seed=1)
eng = Engine(cfg, None)
eng.prepare(); eng.enqueue(30); eng.finish()
print("ok", float(eng.t_timeset[10:30].mean()) * 1e6, "us/round")
eng.close()
