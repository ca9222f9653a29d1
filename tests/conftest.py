import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden", "serra09_golden.npz")
# the CPU tests step the kernel sources under the SIMT emulator (tests/emu/): acoss_b200._native refuses any library
# but the shipped CUDA one without this opt-in (inherited by the torchrun subprocesses of test_parallel.py)
os.environ.setdefault("ACOSS_B200_ALLOW_TEST_LIB", "1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    return np.load(GOLDEN, allow_pickle=False)


def golden_case_songs(G, tag):
    lengths = G["%s_lengths" % tag]
    songs = []
    for si in range(len(lengths)):
        songs.append({"chroma": G["%s_chroma_%d" % (tag, si)], "gchroma": G["%s_gchroma_%d" % (tag, si)]})
    return songs


def golden_cases(G):
    return [dict(zip(G["cases"].dtype.names, row)) for row in G["cases"].tolist()]
