"""Songs with exactly repeated frames (silence, a sustained frame): what the product's "strictly below the separator"
rule costs against the reference's "exactly k, ties by rounding noise" (DESIGN.md section 2 deviation 1).
The reference side is tests/golden/ties_golden.npz (oracle/make_golden_ties.py, the reference's own CRPUtils.py and
SequenceAlignment.c).  The CPU test steps the kernel sources under the SIMT emulator, the GPU test calls the real
library; both print the measured mismatch, which DESIGN.md quotes."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "emu"))

import parity_checks as pc  # noqa: E402
from acoss_b200 import _native  # noqa: E402


def _golden():
    return np.load(os.path.join(HERE, "golden", "ties_golden.npz"))


def _report(st, where):
    q, d = np.array(st["qmax_rel"]), np.array(st["dmax_rel"])
    print("\n[ties/%s] %d pairs, %d with tied lines; on those: %d of %d reference bits missing, %d extra, "
          "%d cells on tied lines of %d; scores equal on %d pairs, |dQmax| median %.3f max %.3f, "
          "|dDmax| median %.3f max %.3f (relative)" % (
              where, st["pairs"], st["pairs_tied"], st["bits_missing"], st["ones_ref"], st["bits_extra"],
              st["cells_on_tied_lines"], st["cells"], st["pairs_score_equal"],
              float(np.median(q)), float(q.max()), float(np.median(d)), float(d.max())))


def _check(st, pairs=18, tied=10, equal=4):
    assert st["pairs"] == pairs and st["pairs_tied"] >= tied
    # tied lines lose their tied cells (fewer than k neighbours); cells the reference lacks appear only where its
    # near-ties (below 1e-9 relative) order differently in the product's arithmetic
    assert st["bits_missing"] <= 0.2 * st["ones_ref"]
    assert st["bits_extra"] <= 0.02 * st["ones_ref"]
    assert st["pairs_score_equal"] >= equal


def test_ties_emulator():
    from build_emu import build
    ctx = _native.Context(0, build())
    try:
        st = pc.check_ties_golden(ctx, _golden(), pairs_per_case=2)   # six pairs here, all eighteen on the GPU
    finally:
        ctx.close()
    _report(st, "emulator")
    _check(st, pairs=6, tied=3, equal=1)


@pytest.mark.gpu
def test_ties_gpu():
    ctx = _native.Context(0)
    try:
        st = pc.check_ties_golden(ctx, _golden())
    finally:
        ctx.close()
    _report(st, "b200")
    _check(st)
