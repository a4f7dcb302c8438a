#!/usr/bin/env python
"""Run the UNMODIFIED reference sources under Python 3 with Python-2 division semantics.

The reference (``baseline/_ref``, a byte-identical copy of /root/reference made by
``install_ref.sh``) is Python-2 code: it has no ``from __future__ import division`` and uses
``int / int`` as an index in a dozen places, which is floor division there and a
float under Python 3.  Instead of editing the files, this launcher compiles every module that
is imported from the reference tree through an AST pass that rewrites ``a / b`` (and ``a /= b``)
into a call of ``__py2_div__``, which floors when both operands are integers (Python ints, numpy
integer scalars or integer arrays) and is true division otherwise -- exactly the "classic
division" the files were written for.  Nothing else is touched: same drivers, same numpy / scipy
/ sklearn calls, same file formats, same timers.

    python baseline/py2run.py main.py <13 positionals>          (cwd = baseline/_ref)

``EH_NP_SEED=<int>`` seeds numpy's global generator before the script starts.  The reference draws its
initial parameters (naive / replication / approximate) and its code matrix B from that generator without
seeding it; with the variable set the run is reproducible, which is what the parity tests compare against.
The benchmark arm leaves it unset.

``EH_REF_API_SHIMS=1`` restores two library interfaces the reference's ``arrange_real_data.py`` was written
against and that current pandas / scikit-learn no longer have: ``DataFrame.ix`` (label-based here,
which is how the script uses it, arrange_real_data.py:38) and ``OneHotEncoder(sparse=...)`` (now
``sparse_output``, :61,111,175,218).  Like the division pass this changes the environment, not the file.
"""
import ast
import builtins
import importlib.abc
import importlib.machinery
import importlib.util
import os
import sys

import numpy as np


def _is_int(x):
    if isinstance(x, bool):
        return True
    if isinstance(x, (int, np.integer)):
        return True
    return isinstance(x, np.ndarray) and x.dtype.kind in "iub"


def __py2_div__(a, b):
    if _is_int(a) and _is_int(b):
        return a // b
    return a / b


builtins.__py2_div__ = __py2_div__


class _ClassicDivision(ast.NodeTransformer):
    def visit_BinOp(self, node):
        self.generic_visit(node)
        if isinstance(node.op, ast.Div):
            call = ast.Call(func=ast.Name(id="__py2_div__", ctx=ast.Load()), args=[node.left, node.right], keywords=[])
            return ast.copy_location(call, node)
        return node

    def visit_AugAssign(self, node):
        self.generic_visit(node)
        if isinstance(node.op, ast.Div) and isinstance(node.target, ast.Name):
            load = ast.Name(id=node.target.id, ctx=ast.Load())
            call = ast.Call(func=ast.Name(id="__py2_div__", ctx=ast.Load()), args=[load, node.value], keywords=[])
            return ast.copy_location(ast.Assign(targets=[node.target], value=call), node)
        return node


def compile_py2(source, filename):
    tree = _ClassicDivision().visit(ast.parse(source, filename))
    ast.fix_missing_locations(tree)
    return compile(tree, filename, "exec", dont_inherit=True)


class _Loader(importlib.machinery.SourceFileLoader):
    def source_to_code(self, data, path, *, _optimize=-1):
        return compile_py2(data, path)

    def get_code(self, fullname):          # never use / write .pyc for the transformed modules
        return self.source_to_code(self.get_data(self.path), self.path)


class _Finder(importlib.abc.MetaPathFinder):
    def __init__(self, roots):
        self.roots = roots

    def find_spec(self, fullname, path=None, target=None):
        if "." in fullname:
            return None
        for r in self.roots:
            f = os.path.join(r, fullname + ".py")
            if os.path.isfile(f):
                return importlib.util.spec_from_file_location(fullname, f, loader=_Loader(fullname, f))
        return None


def _install_api_shims():
    import pandas as pd
    from sklearn import preprocessing
    if not hasattr(pd.DataFrame, "ix"):
        pd.DataFrame.ix = property(lambda self: self.loc)
    base = preprocessing.OneHotEncoder

    class OneHotEncoder(base):
        def __init__(self, *, sparse=None, **kw):
            if sparse is not None:
                kw.setdefault("sparse_output", bool(sparse))
            super().__init__(**kw)

        @classmethod
        def _get_param_names(cls):
            return base._get_param_names()
    preprocessing.OneHotEncoder = OneHotEncoder


def main():
    if len(sys.argv) < 2:
        print(__doc__)
        return 2
    script = os.path.abspath(sys.argv[1])
    root = os.path.dirname(script)
    sys.meta_path.insert(0, _Finder([os.path.join(root, "src"), root]))
    sys.argv = [sys.argv[1]] + sys.argv[2:]
    sys.path.insert(0, root)
    if os.environ.get("EH_REF_API_SHIMS"):
        _install_api_shims()
    if os.environ.get("EH_NP_SEED"):
        np.random.seed(int(os.environ["EH_NP_SEED"]))
    with open(script, "rb") as f:
        code = compile_py2(f.read(), script)
    g = {"__name__": "__main__", "__file__": script, "__builtins__": builtins}
    exec(code, g)
    return 0


if __name__ == "__main__":
    sys.exit(main())
