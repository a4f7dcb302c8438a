#!/bin/bash
# Baseline B1: a *patched* copy of the reference under baseline/_ref_py3 (git-ignored).
# The official reference arm (baseline/_ref, unmodified) cannot be installed: the reference has no
# setup.py/pyproject, and it is Python-2 + mpi4py code.  This script makes the mechanical py3
# patch of SURVEY Q10 (integer '/' -> '//' at the listed sites) so the reference's own drivers can
# be timed on CPU over the in-repo mpi4py subset (baseline/mpi_shim).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
SRC="${1:-/root/reference}"
DST="$HERE/_ref_py3"
if [ ! -f "$SRC/main.py" ]; then
  # e.g. on a GPU box, where only the repository travels: keep an existing patched copy
  [ -f "$DST/main.py" ] && { echo "no reference at $SRC; keeping $DST"; exit 0; }
  echo "no reference at $SRC and no patched copy at $DST" >&2; exit 1
fi
rm -rf "$DST"; mkdir -p "$DST"
cp -r "$SRC/main.py" "$SRC/src" "$DST/"
cd "$DST"
sed -i 's#n_samples/(n_procs-1)#n_samples//(n_procs-1)#g; s#n_workers/(n_stragglers+1)#n_workers//(n_stragglers+1)#g; s#(src-1)/(n_stragglers+1)#(src-1)//(n_stragglers+1)#g; s#(rank-1)/(n_stragglers+1)#(rank-1)//(n_stragglers+1)#g; s#n_samples/((n_partitions-n_stragglers)\*n_workers)#n_samples//((n_partitions-n_stragglers)*n_workers)#g' src/*.py
sed -i 's#n_rows_per_worker = n_rows / partitions#n_rows_per_worker = n_rows // partitions#; s#idx=(st+nd)/2#idx=(st+nd)//2#' src/generate_data.py src/util.py
echo "patched reference in $DST"
