"""CPU-only tests of the host side: C-ABI surface, HDF5 io, the plugin mirror's pair enumeration,
batch blocks, evaluation statistics and feature aggregation."""
import itertools
import os
import sys
import re

import numpy as np
import pytest

from acoss_b200 import _native, hdf5io, synth
from acoss_b200.cover_algorithm import CoverAlgorithm
from acoss_b200.serra09 import Serra09, global_chroma, sync_aggregate

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "acoss_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(acoss_[a-z0-9_]+)\s*\(", text)))


def test_abi_exports_every_declared_symbol():
    if not os.path.exists(_native.DEFAULT_LIB):
        from acoss_b200 import build
        build.build()
    lib = _native.load_library()  # binds every name in SIGNATURES or raises
    names = declared_symbols()
    assert len(names) >= 16
    for n in names:
        assert hasattr(lib, n), n
        assert n in _native.SIGNATURES, "binding missing for %s" % n
    assert sorted(_native.SIGNATURES) == names
    assert lib.acoss_version() >= 100


def test_no_gpu_means_loud_failure_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(_native.AcossError):
        _native.Context(0)


def test_hdf5_roundtrip(tmp_path):
    d = {"chroma_qmax": np.random.rand(7, 7).astype(np.float32),
         "similarities": {"idxs": np.arange(10).reshape(5, 2), "chroma_dmax": np.random.rand(5),
                          "empty": np.zeros((0, 2), np.int64)},
         "label": "clique_1", "n": 3, "f": 2.5, "flag": True, "nothing": None,
         "hpcp": np.random.rand(20, 12).astype(np.float32), "u8": np.arange(5, dtype=np.uint8)}
    p = str(tmp_path / "t.h5")
    hdf5io.save(p, d)
    r = hdf5io.load(p)
    assert sorted(r) == sorted(d)
    for k in ("chroma_qmax", "hpcp", "u8"):
        assert r[k].dtype == d[k].dtype and np.array_equal(r[k], d[k])
    assert np.array_equal(r["similarities"]["idxs"], d["similarities"]["idxs"])
    assert r["similarities"]["empty"].shape == (0, 2)
    assert (r["label"], r["n"], r["f"], r["nothing"]) == ("clique_1", 3, 2.5, None)
    # many entries in one group, and the file starts with the HDF5 signature + superblock v0
    big = {"k%03d" % i: np.full(3, i) for i in range(150)}
    hdf5io.save(p, big)
    r = hdf5io.load(p)
    assert len(r) == 150 and r["k042"][0] == 42
    raw = open(p, "rb").read(16)
    assert raw[:8] == b"\x89HDF\r\n\x1a\n" and raw[8] == 0


def test_all_pairs_order_matches_itertools():
    songs, labels = synth.make_song_set("covers80", seed=1, n_songs=7, length_range=(20, 30))
    alg = CoverAlgorithm(songs=songs, labels=labels, do_memmaps=False)
    assert [tuple(p) for p in alg.all_pairs_list(True)] == list(itertools.combinations(range(7), 2))
    assert [tuple(p) for p in alg.all_pairs_list(False)] == list(itertools.permutations(range(7), 2))


def test_batch_blocks_cover_lower_triangle():
    # runcovers80.sh: w = 20 over 160 songs -> 36 blocks (CoverAlgorithm.py:227-244)
    songs = [{"chroma": np.zeros((12, 1), np.float32), "gchroma": np.ones(12)} for _ in range(40)]
    alg = CoverAlgorithm(songs=songs, labels=["x"] * 40, do_memmaps=False)
    w, wsub = 10, 5
    seen = set()
    nblocks = (40 // w) * (40 // w + 1) // 2
    for idx in range(nblocks):
        for i in range(w // wsub):
            for j in range(w // wsub):
                for a, b in alg.batch_block_idxs(w, idx, wsub, i, j):
                    assert a >= b
                    assert (a, b) not in seen
                    seen.add((int(a), int(b)))
    assert seen == {(a, b) for a in range(40) for b in range(a + 1)}


def reference_eval(D, cliques):
    """Direct restatement of the rank loop of CoverAlgorithm.py:345-395 (the O(N^2) version)."""
    from itertools import chain
    N = D.shape[0]
    cl = [list(c) for c in cliques]
    Ks = np.array([len(c) for c in cl])
    idx = np.argsort(-Ks, kind="stable")
    Ks = Ks[idx]
    cl = [cl[i] for i in idx]
    idx = np.array(list(chain(*cl)), dtype=int)
    D = D[idx, :][:, idx].copy()
    np.fill_diagonal(D, -np.inf)
    order = np.argsort(-D, 1)
    ranks = np.nan * np.ones(N)
    AllMap = np.nan * np.ones(N)
    startidx, kidx = 0, 0
    for i in range(N):
        if i >= startidx + Ks[kidx]:
            startidx += Ks[kidx]
            kidx += 1
            if Ks[kidx] < 2:
                break
        iranks = [k + 1 for k in range(N) if 0 <= order[i, k] - startidx < Ks[kidx]][0:-1]
        ranks[i] = iranks[0]
        AllMap[i] = np.mean([float(j) / float(r) for j, r in zip(range(1, Ks[kidx]), iranks)])
    ranks = ranks[~np.isnan(ranks)]
    return np.mean(ranks), np.sum(1.0 / ranks) / N, np.median(ranks), np.nanmean(AllMap)


def test_eval_statistics_match_reference_loop(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    rng = np.random.default_rng(3)
    labels = ["a"] * 3 + ["b"] * 2 + ["c"] * 2 + ["d", "e", "f"]
    perm = rng.permutation(len(labels))
    labels = [labels[p] for p in perm]
    n = len(labels)
    songs = [{"chroma": np.zeros((12, 1), np.float32), "gchroma": np.ones(12), "label": labels[i]} for i in range(n)]
    alg = CoverAlgorithm(songs=songs, labels=labels, similarity_types=["main"], do_memmaps=False, cachedir=str(tmp_path))
    D = rng.uniform(size=(n, n)).astype(np.float32)
    D = D + D.T
    alg.Ds["main"] = D
    alg.get_all_clique_ids()
    # the rank loop runs on the device (acoss_eval_ranks); without a GPU the same kernel source is stepped by the
    # SIMT emulator (test infrastructure) so that the host half of getEvalStatistics is still covered here
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))
    from build_emu import build
    from acoss_b200 import _native
    alg._eval_ctx = _native.Context(0, build())
    MR, MRR, MDR, MAP, tops = alg.getEvalStatistics("main", topsidx=[1, 3])
    alg._eval_ctx.close()
    ref = reference_eval(D, list(alg.cliques.values()))
    assert np.allclose([MR, MRR, MDR, MAP], ref)
    assert os.path.exists("results_full.csv")
    assert open("results_full.csv").readline().startswith("name, MR, MRR, MDR, MAP,Top-1")


def test_feature_aggregation_matches_definition():
    rng = np.random.default_rng(0)
    chroma = rng.uniform(size=(95, 12)).astype(np.float32)   # (n_frames, 12) like hpcp
    g = global_chroma(chroma)
    assert g.shape == (12,) and g.max() == 1.0
    agg = sync_aggregate(chroma.T, 40, np.median)
    assert agg.shape == (12, 3)
    assert np.array_equal(agg[:, 0], np.median(chroma[:40].T, axis=1))
    assert np.array_equal(agg[:, 2], np.median(chroma[80:].T, axis=1))
    with pytest.raises(IOError):
        global_chroma(np.zeros((5, 13)))


def test_serra09_load_features_from_files(tmp_path):
    # per-song .h5 with the fields extractors.py writes (hpcp, mfcc_htk, label)
    rng = np.random.default_rng(1)
    for s in range(3):
        hdf5io.save(str(tmp_path / ("song%d.h5" % s)),
                    {"hpcp": rng.uniform(size=(400 + 37 * s, 12)).astype(np.float32),
                     "mfcc_htk": rng.normal(size=(13, 400 + 37 * s)).astype(np.float32),
                     "label": "c%d" % (s % 2), "track_id": "t%d" % s})
    alg = Serra09(datapath=str(tmp_path), chroma_type="hpcp", shortname="t", do_memmaps=False)
    assert alg.N == 3
    f = alg.load_features(2)
    L = int(np.ceil((400 + 74) / 40))
    assert f["chroma"].shape == (12, L) and f["mfcc"].shape == (13, L) and f["gchroma"].shape == (12,)
    assert alg.cliques == {"c0": {2}}


def test_hdf5_chunked_gzip_shuffle_roundtrip_and_filters(tmp_path):
    """Chunked datasets with the shuffle + gzip filters (what deepdish / PyTables write with compression='zlib'):
    several chunks, ragged last chunk, 1-D / 2-D / 3-D, every numeric kind; the filter-pipeline parser on version-1 and
    version-2 messages; blosc is refused with an instruction to re-export.  No HDF5 library exists here, so the file is
    assembled by hdf5io's own writer from the format specification (interop unpinned, see hdf5io's docstring)."""
    import struct
    rng = np.random.default_rng(0)
    d = {"hpcp": rng.random((30011, 12)).astype(np.float32), "crema": rng.random((7001, 12)),
         "mfcc_htk": (rng.standard_normal((13, 5000)) * 30).astype(np.float32), "ids": np.arange(300000, dtype=np.int64),
         "cube": rng.integers(0, 255, size=(70, 50, 40)).astype(np.uint8), "label": "W_123", "track_id": "P_1",
         "madmom_features": {"onsets": rng.random(77), "tempos": rng.random((3, 2)), "empty": np.zeros((0, 2))}}
    path = str(tmp_path / "song.h5")
    hdf5io.save(path, d, compression="zlib")
    plain = str(tmp_path / "plain.h5")
    hdf5io.save(plain, d)
    assert os.path.getsize(path) < os.path.getsize(plain)
    out = hdf5io.load(path)
    for k in ("hpcp", "crema", "mfcc_htk", "ids", "cube"):
        assert out[k].dtype == d[k].dtype and np.array_equal(out[k], d[k]), k
    assert out["label"] == "W_123" and np.array_equal(out["madmom_features"]["tempos"], d["madmom_features"]["tempos"])
    assert out["madmom_features"]["empty"].shape == (0, 2)
    # filter pipeline messages: version 1 (padded names) and version 2 (no name for ids < 256)
    v1 = struct.pack("<BB6x", 1, 2) + struct.pack("<HHHH", 2, 8, 0, 1) + b"shuffle\0" + struct.pack("<II", 4, 0) \
        + struct.pack("<HHHH", 1, 8, 1, 1) + b"deflate\0" + struct.pack("<II", 5, 0)
    v2 = struct.pack("<BB", 2, 2) + struct.pack("<HHH", 2, 0, 1) + struct.pack("<I", 4) + struct.pack("<HHH", 1, 1, 1) + struct.pack("<I", 5)
    assert hdf5io._Reader.parse_filters(v1) == [2, 1] and hdf5io._Reader.parse_filters(v2) == [2, 1]
    blosc = struct.pack("<BB", 2, 1) + struct.pack("<HHHH", 32001, 6, 0, 0) + b"blosc\0"
    assert hdf5io._Reader.parse_filters(blosc) == [32001]
    r = hdf5io._Reader(np.memmap(path, dtype=np.uint8, mode="r"))
    with pytest.raises(NotImplementedError, match="blosc.*re-export"):
        r.read_chunked(struct.pack("<BBB", 3, 2, 2) + struct.pack("<Q", 0) + struct.pack("<II", 4, 4), np.dtype("f4"), (8,), [32001])


def test_only_the_shipped_library_loads_without_opt_in(monkeypatch, tmp_path):
    """acoss_b200._native refuses a library other than the in-tree CUDA build unless the test opt-in is set: the
    emulator build can never be picked up as a fallback."""
    from acoss_b200 import _native
    monkeypatch.delenv("ACOSS_B200_ALLOW_TEST_LIB", raising=False)
    other = tmp_path / "libother.so"
    other.write_bytes(b"")
    with pytest.raises(_native.AcossError, match="refusing to load"):
        _native.load_library(str(other))


def test_bench_reference_arm_prints_one_json_line():
    """bench.py --impl reference (the CPU arm the driver runs beside the GPU arm): one JSON line on stdout with the
    contract's keys, on a tiny sample so that it runs in seconds here."""
    import json
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--config", "covers80",
                        "--cpu-pairs", "24", "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "song_pairs_per_sec" and d["unit"] == "pairs/s"
    assert d["value"] > 0 and d["higher_is_better"] is True and d["steps"] == 1
    assert d["cpu_baseline"]["kind"] in ("port", "reference") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"] == {"value": d["value"], "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["songs"] == 160 and d["config"]["m"] == 9
