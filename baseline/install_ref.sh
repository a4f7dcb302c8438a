#!/bin/bash
# Install the reference for `bench.py --impl reference` (baseline B1) under baseline/_ref.
#
# The prescribed `pip install --no-index --no-build-isolation --find-links /opt/wheelhouse
# --target baseline/_ref /root/reference` fails ("Neither 'setup.py' nor 'pyproject.toml' found":
# the reference is a directory of scripts, not a package), so the tree is copied VERBATIM
# (byte-identical; checked below with cmp).  It is Python-2 + mpi4py code; baseline/py2run.py
# gives it Python-2 division semantics at import time and baseline/mpi_shim provides the mpi4py
# calls it uses, so no file of the reference is edited.  baseline/_ref is git-ignored and ships
# to the GPU box with the repository snapshot.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
SRC="${1:-/root/reference}"
DST="$HERE/_ref"
if [ ! -f "$SRC/main.py" ]; then
  # e.g. on a GPU box, where only the repository travels: keep an existing copy
  [ -f "$DST/main.py" ] && { echo "no reference at $SRC; keeping $DST"; exit 0; }
  echo "no reference at $SRC and no copy at $DST" >&2; exit 1
fi
python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target "$DST" "$SRC" \
  > "$HERE/.pip_install.log" 2>&1 || echo "pip install of the reference failed (expected, see header): $(tail -1 "$HERE/.pip_install.log")"
rm -rf "$DST"; mkdir -p "$DST"
cp -r "$SRC/main.py" "$SRC/src" "$DST/"
for f in README.md run_approx_coding.sh data_prepare.sh Makefile Makefile.make; do
  [ -f "$SRC/$f" ] && cp "$SRC/$f" "$DST/"
done
for f in $(cd "$SRC" && ls main.py src/*.py); do cmp -s "$SRC/$f" "$DST/$f" || { echo "copy of $f differs" >&2; exit 1; }; done
rm -f "$HERE/.pip_install.log"
echo "reference copied unmodified to $DST"
