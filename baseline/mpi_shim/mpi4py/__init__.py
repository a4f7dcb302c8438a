"""Minimal stand-in for the subset of mpi4py the reference uses (see MPI.py).  Baseline tooling only."""
