"""acoss_dp_packed_static (csrc/noalloc.cu): the DPs with their scratch in a static __device__ array, for callers that
cannot allocate.  The new entry point is called through the C ABI, twice in a row, and compared with the existing
acoss_dp_packed on the same plots: Qmax / Dmax exact (half-unit integers), Smith-Waterman within the 1e-5 relative
tolerance of tests/parity_checks.py.  The second call uses fewer, differently shaped plots, so anything the shared
array kept from the first call that should have been rewritten would show.  In a file of its own that pytest collects
last."""
import ctypes

import numpy as np
import pytest

from acoss_b200 import _native

pytestmark = pytest.mark.gpu

SW_RTOL = 1e-5


def _plots(seed, shapes):
    rng = np.random.default_rng(seed)
    return [(rng.uniform(size=s) < dens).astype(np.uint8) for s, dens in shapes]


def _run_static(lib, device, plots, flags):
    import torch
    packed = [_native.pack_bits(p) for p in plots]
    words = np.array([p.size for p in packed], np.int64)
    offs = np.zeros(len(plots), np.int64)
    if len(plots) > 1:
        offs[1:] = np.cumsum(words[:-1])
    flat = np.concatenate([p.reshape(-1) for p in packed]).astype(np.uint32)
    # the CALLER owns inputs and outputs: allocated here, before the call, not by the entry point
    crp = torch.from_numpy(flat.view(np.int32)).to("cuda:%d" % device)
    out = torch.full((3 * len(plots),), -7.0, dtype=torch.float32, device="cuda:%d" % device)
    torch.cuda.synchronize(device)
    _native.dp_packed_static(lib, device, crp.data_ptr(), offs, [p.shape[0] for p in plots], [p.shape[1] for p in plots],
                             out.data_ptr(), flags=flags, stream=0)
    torch.cuda.synchronize(device)
    return out.cpu().numpy().reshape(3, len(plots)).T


def _check(got, ref, want_sw, where):
    assert np.array_equal(got[:, 0], ref[:, 0]), where + " qmax"
    assert np.array_equal(got[:, 1], ref[:, 1]), where + " dmax"
    if want_sw:
        assert np.all(np.abs(got[:, 2] - ref[:, 2]) <= SW_RTOL * np.maximum(1.0, np.abs(ref[:, 2]))), where + " sw"
    else:
        assert not got[:, 2].any(), where + " sw must be zero without ACOSS_FLAG_WANT_SW"


def test_static_scratch_entry_matches_dp_packed_twice_in_a_row():
    ctx = _native.Context(0)  # raises when the CUDA library or the GPU is missing
    try:
        lib = ctx.lib
        # call 1: 40 plots, narrow and wide (several 128-column blocks -> hand-over records in the static scratch)
        shapes1 = [((40, c), d) for c in (4, 33, 127, 128, 129, 257, 385, 700) for d in (0.08, 0.5)]
        shapes1 += [((r, 300), 0.3) for r in (3, 4, 5, 130, 600)] + [((1200, 1100), 0.06), ((2, 2), 0.5), ((0, 5), 0.5)]
        shapes1 += [((64 + 7 * t, 90 + 31 * t), 0.1) for t in range(16)]
        plots1 = _plots(1, shapes1)
        assert len(plots1) == 40
        # call 2: fewer plots, other shapes, denser
        shapes2 = [((77, 515), 0.4), ((300, 140), 0.2), ((5, 6), 0.9), ((900, 131), 0.07), ((33, 33), 0.5), ((450, 460), 0.055)]
        plots2 = _plots(2, shapes2)
        for flags, want_sw in ((0, False), (_native.FLAG_DMAX_CLEAN | _native.FLAG_WANT_SW, True)):
            ref1 = ctx.dp_packed(plots1, flags=flags & ~_native.FLAG_WANT_SW, want_sw=want_sw)
            ref2 = ctx.dp_packed(plots2, flags=flags & ~_native.FLAG_WANT_SW, want_sw=want_sw)
            got1 = _run_static(lib, 0, plots1, flags)
            got2 = _run_static(lib, 0, plots2, flags)
            got1b = _run_static(lib, 0, plots1, flags)
            _check(got1, ref1, want_sw, "first call, flags %d" % flags)
            _check(got2, ref2, want_sw, "second call, flags %d" % flags)
            _check(got1b, ref1, want_sw, "third call, flags %d" % flags)
    finally:
        ctx.close()


def test_static_scratch_entry_rejects_sizes_beyond_its_maximum():
    ctx = _native.Context(0)
    try:
        lib = ctx.lib
        import torch
        crp = torch.zeros(64, dtype=torch.int32, device="cuda:0")
        out = torch.zeros(3 * 80, dtype=torch.float32, device="cuda:0")
        one = np.zeros(1, np.int64)
        with pytest.raises(_native.AcossError, match="at most 64"):
            _native.dp_packed_static(lib, 0, crp.data_ptr(), np.zeros(65, np.int64), np.ones(65, np.int32), np.ones(65, np.int32),
                                     out.data_ptr())
        with pytest.raises(_native.AcossError, match="limit is 16384"):
            _native.dp_packed_static(lib, 0, crp.data_ptr(), one, [20000], [4], out.data_ptr())
        with pytest.raises(_native.AcossError, match="holds 65536"):
            _native.dp_packed_static(lib, 0, crp.data_ptr(), np.zeros(5, np.int64), [16000] * 5, [200] * 5, out.data_ptr())
        assert lib.acoss_dp_packed_static(0, None, None, None, None, 0, 0, None, None) == 0   # B = 0: nothing to do
    finally:
        ctx.close()
