"""oracle/make_golden_eval.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Generates tests/golden/eval_golden.npz by running THE REFERENCE'S OWN getEvalStatistics
(/root/reference/benchmarking/CoverAlgorithm.py:330-418).  The module cannot be imported
(deepdish / matplotlib at module level), so the function's source text is cut out
of the file with `ast` and executed unchanged against a stub object that carries the four
attributes it reads (Ds, cliques, name, shortname).  Nothing of the reference is copied into the
repo: only the inputs and the numbers it returned.
Run:  python -m oracle.make_golden_eval      (needs /root/reference)
"""
from __future__ import annotations

import ast
import contextlib
import io
import os
import sys
import tempfile
import textwrap
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

REF = "/root/reference/benchmarking/CoverAlgorithm.py"


def load_reference_eval():
    src = open(REF).read()
    tree = ast.parse(src)
    for node in ast.walk(tree):
        if isinstance(node, ast.FunctionDef) and node.name == "getEvalStatistics":
            lines = src.splitlines()[node.lineno - 1:node.end_lineno]
            ns = {"np": np, "warnings": warnings, "os": os}
            exec(textwrap.dedent("\n".join(lines)), ns)
            return ns["getEvalStatistics"]
    raise RuntimeError("getEvalStatistics not found in %s" % REF)


class Stub(object):
    def __init__(self, D, cliques):
        self.Ds = {"main": D}
        self.cliques = {"clique_%i" % i: set(c) for i, c in enumerate(cliques)}
        self.name = "golden"
        self.shortname = "golden"


def make_case(rng, sizes, boost=0.6, tie_free=True):
    """Score matrix with cover structure: clique members score higher on average."""
    n = int(np.sum(sizes))
    perm = rng.permutation(n)
    cliques, at = [], 0
    for K in sizes:
        cliques.append([int(x) for x in perm[at:at + K]])
        at += K
    label = np.zeros(n, int)
    for c, members in enumerate(cliques):
        label[members] = c
    D = rng.random((n, n)).astype(np.float32) * 0.5
    same = label[:, None] == label[None, :]
    D[same] += (rng.random(int(same.sum())) * boost).astype(np.float32)
    D = np.triu(D, 1)
    D = D + D.T
    if tie_free:
        # distinct values within every row (the reference's introsort is only defined then)
        for i in range(n):
            assert np.unique(np.delete(D[i], i)).size == n - 1
    return D.astype(np.float32), cliques


def main():
    fn = load_reference_eval()
    rng = np.random.default_rng(330)
    out = {}
    cases = [("a", [2] * 20, 0.6), ("b", [13] * 3 + [1] * 11, 0.2), ("c", [5, 4, 3, 2, 2, 1, 1, 1], 0.3),
             ("d", [40, 2, 1], 0.05), ("e", [1] * 6 + [3] * 4 + [2] * 5, 0.4), ("f", [1] * 9, 0.5)]
    tmp = tempfile.mkdtemp()
    cwd = os.getcwd()
    os.chdir(tmp)   # the function appends to results_<shortname>.csv in the working directory
    try:
        for tag, sizes, boost in cases:
            D, cliques = make_case(rng, sizes, boost)
            # the reference keeps cliques as sets (CoverAlgorithm.py:92-106) and unrolls them with list(set):
            # store the songs in that iteration order, it fixes the summation order of MAP
            cliques = [list(set(c)) for c in cliques]
            stub = Stub(D, cliques)
            with contextlib.redirect_stdout(io.StringIO()), warnings.catch_warnings():
                warnings.simplefilter("ignore")
                MR, MRR, MDR, MAP, tops = fn(stub, "main", topsidx=[1, 10, 100, 1000])
            out["%s_D" % tag] = D
            out["%s_clique_sizes" % tag] = np.array([len(c) for c in cliques], np.int32)
            out["%s_clique_songs" % tag] = np.array([s for c in cliques for s in c], np.int32)
            out["%s_stats" % tag] = np.array([MR, MRR, MDR, MAP], np.float64)
            out["%s_tops" % tag] = np.asarray(tops, np.float64)
            print(tag, sizes, "->", MR, MRR, MDR, MAP, tops)
    finally:
        os.chdir(cwd)
    out["tags"] = np.array([c[0] for c in cases])
    path = os.path.join(ROOT, "tests", "golden", "eval_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path)


if __name__ == "__main__":
    main()
