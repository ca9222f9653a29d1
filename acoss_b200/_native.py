"""ctypes binding of libacoss_b200.so (the C ABI in include/acoss_b200.h).

This is the only place the package talks to native code.  There is NO fallback: if the
CUDA library has not been built (``python -m acoss_b200.build``) or there is no usable
GPU, importing works but creating a context raises.  PyTorch is not needed here; numpy
arrays are handed over as host pointers.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_LIB = os.path.join(HERE, "lib", "libacoss_b200.so")

F32, F64 = 0, 1
FLAG_NO_OTI = 1 << 0
FLAG_DMAX_CLEAN = 1 << 1
FLAG_NO_SWAP = 1 << 2
FLAG_NON_MUTUAL = 1 << 3
FLAG_WANT_SW = 1 << 4
FLAG_DEVICE_INPUT = 1 << 8

_c_i32p = ctypes.POINTER(ctypes.c_int32)
_c_i64p = ctypes.POINTER(ctypes.c_int64)
_c_u32p = ctypes.POINTER(ctypes.c_uint32)
_c_u8p = ctypes.POINTER(ctypes.c_uint8)
_c_f32p = ctypes.POINTER(ctypes.c_float)
_c_f64p = ctypes.POINTER(ctypes.c_double)

# name -> (restype, argtypes); every symbol include/acoss_b200.h declares
SIGNATURES = {
    "acoss_last_error": (ctypes.c_char_p, []),
    "acoss_version": (ctypes.c_int, []),
    "acoss_create": (ctypes.c_int, [ctypes.POINTER(ctypes.c_void_p), ctypes.c_int]),
    "acoss_destroy": (ctypes.c_int, [ctypes.c_void_p]),
    "acoss_upload_songs": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, _c_i64p, _c_i32p,
                                          ctypes.c_int32, ctypes.c_int32, _c_f64p]),
    "acoss_serra09_pairs": (ctypes.c_int, [ctypes.c_void_p, _c_i32p, ctypes.c_int64, ctypes.c_int32, ctypes.c_double,
                                           ctypes.c_uint32, _c_f32p, _c_f32p, _c_f32p, _c_i32p, _c_i32p, _c_u32p,
                                           _c_i64p, _c_i32p]),
    "acoss_align_dense": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, _c_i64p, _c_i32p, _c_i32p,
                                         ctypes.c_int32, ctypes.c_double, ctypes.c_uint32, _c_f32p, _c_u32p, _c_i64p]),
    "acoss_aggregate_songs": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, _c_i64p, _c_i32p,
                                             ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                             ctypes.c_void_p, _c_i64p, ctypes.c_void_p]),
    "acoss_eval_ranks": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, _c_i32p, _c_i32p, _c_i32p,
                                        ctypes.c_int32, _c_i64p, _c_i32p, ctypes.c_uint32]),
    "acoss_last_eval_ms": (ctypes.c_int, [ctypes.c_void_p, _c_f32p]),
    "acoss_dp_packed": (ctypes.c_int, [ctypes.c_void_p, _c_u32p, _c_i64p, _c_i32p, _c_i32p, ctypes.c_int32,
                                       ctypes.c_uint32, _c_f32p]),
    "acoss_qmax": (ctypes.c_int, [ctypes.c_void_p, _c_u8p, ctypes.c_int32, ctypes.c_int32, _c_f32p]),
    "acoss_dmax": (ctypes.c_int, [ctypes.c_void_p, _c_u8p, ctypes.c_int32, ctypes.c_int32, _c_f32p]),
    "acoss_qmax_dmax": (ctypes.c_int, [ctypes.c_void_p, _c_u8p, ctypes.c_int32, ctypes.c_int32, _c_f32p, _c_f32p]),
    "acoss_swconstrained": (ctypes.c_int, [ctypes.c_void_p, _c_u8p, ctypes.c_int32, ctypes.c_int32, _c_f32p]),
    "acoss_set_option": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int64]),
    "acoss_kernel_launches": (ctypes.c_int64, [ctypes.c_void_p]),
    "acoss_reset_counters": (None, [ctypes.c_void_p]),
    "acoss_bind_score_matrices": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                                                 _c_i32p, ctypes.c_int32]),
    "acoss_last_stage_ms": (ctypes.c_int, [ctypes.c_void_p, _c_f32p]),
    "acoss_debug_host_ms": (ctypes.c_int, [ctypes.c_void_p, _c_f32p]),
    "acoss_last_aggregate_stats": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_double)]),
    "acoss_upload_aux_songs": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int32]),
    "acoss_snf_pairs": (ctypes.c_int, [ctypes.c_void_p, _c_i32p, ctypes.c_int64, ctypes.c_int32, ctypes.c_double, ctypes.c_int32,
                                       ctypes.c_uint32, _c_f32p, _c_f32p, _c_i32p, _c_i32p, _c_f64p, _c_i64p, _c_u32p, _c_i64p]),
    "acoss_dp_packed_static": (ctypes.c_int, [ctypes.c_int32, ctypes.c_void_p, _c_i64p, _c_i32p, _c_i32p, ctypes.c_int32,
                                              ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p]),
    "acoss_debug_sweep7_stats": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint64), ctypes.c_int]),
}


class AcossError(RuntimeError):
    pass


def _ptr(arr, typ):
    return None if arr is None else arr.ctypes.data_as(typ)


_SHIPPED_LIB = DEFAULT_LIB


def load_library(path: str | None = None) -> ctypes.CDLL:
    path = path or DEFAULT_LIB
    # Only the CUDA library built in-tree is the product.  tests/ also load the same kernel sources compiled for the
    # host against a SIMT emulator (tests/emu/); that build is test infrastructure, never a fallback, and loading any
    # library other than the shipped one needs the explicit opt-in that tests/conftest.py sets.
    if os.path.realpath(path) != os.path.realpath(_SHIPPED_LIB) and os.environ.get("ACOSS_B200_ALLOW_TEST_LIB") != "1":
        raise AcossError("refusing to load %s: only %s is the product library (set ACOSS_B200_ALLOW_TEST_LIB=1 in "
                         "tests that load the emulator build)" % (path, _SHIPPED_LIB))
    if not os.path.exists(path):
        raise AcossError("native library %s not found: build it with `python -m acoss_b200.build` "
                         "(there is no CPU fallback)" % path)
    lib = ctypes.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the ABI is incomplete
        fn.restype = res
        fn.argtypes = args
    return lib


def _host_array(shape, dtype, pinned):
    """numpy array, in page-locked host memory when asked (torch is only the allocator here)."""
    if pinned:
        try:
            import torch
            if torch.cuda.is_available():
                t = torch.empty(tuple(shape), dtype=getattr(torch, np.dtype(dtype).name), pin_memory=True)
                return t.numpy()
        except Exception:
            pass
    return np.empty(shape, dtype=dtype)


def pack_songs(songs, pinned=False):
    """songs: list of dicts with 'chroma' (bins, L) [and 'gchroma' (bins,)] as Serra09.load_features
    returns them (Serra09.py:154).  Returns (frames [total, bins], offsets i64, lengths i32,
    gchroma f64 [n, bins] or None)."""
    bins = songs[0]["chroma"].shape[0]
    dt = np.float64 if any(s["chroma"].dtype == np.float64 for s in songs) else np.float32
    lengths = np.array([s["chroma"].shape[1] for s in songs], dtype=np.int32)
    offsets = np.zeros(len(songs), dtype=np.int64)
    if len(songs) > 1:
        offsets[1:] = np.cumsum(lengths[:-1].astype(np.int64))
    total = int(lengths.sum())
    frames = _host_array((max(total, 1), bins), dt, pinned)
    for s, o, n in zip(songs, offsets, lengths):
        frames[o:o + n] = s["chroma"].T
    g = None
    if all(("gchroma" in s and s["gchroma"] is not None) for s in songs):
        g = np.ascontiguousarray(np.stack([np.asarray(s["gchroma"], dtype=np.float64) for s in songs]))
    return frames, offsets, lengths, g


class Context:
    """One acoss_ctx (one GPU)."""

    def __init__(self, device: int = 0, lib_path: str | None = None):
        self.lib = load_library(lib_path)
        self._h = ctypes.c_void_p()
        self.lengths = None
        self.bins = 0
        rc = self.lib.acoss_create(ctypes.byref(self._h), int(device))
        if rc != 0:
            raise AcossError("acoss_create(%d) failed: %s" % (device, self._err()))

    def _err(self):
        e = self.lib.acoss_last_error()
        return e.decode() if e else "?"

    def _check(self, rc, what):
        if rc != 0:
            raise AcossError("%s failed (%d): %s" % (what, rc, self._err()))

    def close(self):
        if self._h:
            self.lib.acoss_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ songs
    def upload(self, frames, offsets, lengths, gchroma=None):
        frames = np.ascontiguousarray(frames)
        if frames.dtype == np.float32:
            dt = F32
        elif frames.dtype == np.float64:
            dt = F64
        else:
            raise TypeError("frames must be float32 or float64, got %s" % frames.dtype)
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        lengths = np.ascontiguousarray(lengths, dtype=np.int32)
        if gchroma is not None:
            gchroma = np.ascontiguousarray(gchroma, dtype=np.float64)
        self.lengths = lengths.copy()
        self.bins = int(frames.shape[1])
        rc = self.lib.acoss_upload_songs(self._h, frames.ctypes.data_as(ctypes.c_void_p), dt,
                                         _ptr(offsets, _c_i64p), _ptr(lengths, _c_i32p), len(lengths),
                                         self.bins, _ptr(gchroma, _c_f64p))
        self._check(rc, "acoss_upload_songs")

    def upload_songs(self, songs):
        self.upload(*pack_songs(songs))

    def upload_device(self, frames_ptr, dtype, bins, offsets, lengths, gchroma_ptr=None):
        """Same as upload() for an arena that already lives in DEVICE memory (addresses as integers, e.g. a torch
        tensor's data_ptr()): frames [total_frames][bins] of float32 / float64, gchroma [n_songs][bins] float64 or
        None.  The copy is device-to-device; the producer of the buffers must have finished."""
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        lengths = np.ascontiguousarray(lengths, dtype=np.int32)
        self.lengths = lengths.copy()
        self.bins = int(bins)
        dt = F32 if np.dtype(dtype) == np.float32 else F64
        rc = self.lib.acoss_upload_songs(self._h, ctypes.c_void_p(int(frames_ptr)), dt, _ptr(offsets, _c_i64p),
                                         _ptr(lengths, _c_i32p), len(lengths), self.bins,
                                         ctypes.cast(ctypes.c_void_p(int(gchroma_ptr)), _c_f64p) if gchroma_ptr else None)
        self._check(rc, "acoss_upload_songs")

    def upload_aux(self, frames):
        """Second feature arena of the songs already uploaded (same offsets / lengths), e.g. mfcc: (total_frames, bins)."""
        frames = np.ascontiguousarray(frames)
        if frames.dtype not in (np.float32, np.float64):
            raise TypeError("frames must be float32 or float64")
        rc = self.lib.acoss_upload_aux_songs(self._h, frames.ctypes.data_as(ctypes.c_void_p),
                                             F32 if frames.dtype == np.float32 else F64, int(frames.shape[1]))
        self._check(rc, "acoss_upload_aux_songs")

    def snf_pairs(self, pairs, m=9, kappa=0.095, niters=3, flags=0, want_fused=False, want_crp=False):
        """EarlySNF's fusion + tail for (P, 2) pairs: dict with qmax_raw, dmax_raw, oti, shape [, fused: list of (M, N)
        float64 = MINUS the fused cross block, crp: list of uint8 (M, N)]."""
        pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
        P = pairs.shape[0]
        q, d = np.zeros(P, np.float32), np.zeros(P, np.float32)
        oti, shape = np.zeros(P, np.int32), np.zeros((P, 2), np.int32)
        L = self.lengths
        Mi = np.maximum(L[pairs[:, 0]].astype(np.int64) - m + 1, 0)
        Nj = np.maximum(L[pairs[:, 1]].astype(np.int64) - m + 1, 0)
        fbuf = foff = cbuf = coff = None
        if want_fused:
            foff = np.concatenate(([0], np.cumsum(Mi * Nj)[:-1])).astype(np.int64)
            fbuf = np.zeros(max(int((Mi * Nj).sum()), 1), np.float64)
        if want_crp:
            words = Mi * ((Nj + 31) // 32)
            coff = np.concatenate(([0], np.cumsum(words)[:-1])).astype(np.int64)
            cbuf = np.zeros(max(int(words.sum()), 1), np.uint32)
        rc = self.lib.acoss_snf_pairs(self._h, _ptr(pairs, _c_i32p), P, int(m), float(kappa), int(niters), int(flags),
                                      _ptr(q, _c_f32p), _ptr(d, _c_f32p), _ptr(oti, _c_i32p), _ptr(shape, _c_i32p),
                                      _ptr(fbuf, _c_f64p), _ptr(foff, _c_i64p), _ptr(cbuf, _c_u32p), _ptr(coff, _c_i64p))
        self._check(rc, "acoss_snf_pairs")
        out = {"qmax_raw": q, "dmax_raw": d, "oti": oti, "shape": shape}
        if want_fused:
            out["fused"] = [fbuf[o:o + a * b].reshape(int(a), int(b)) for o, a, b in zip(foff, Mi, Nj)]
        if want_crp:
            out["crp"] = [unpack_bits(cbuf[o:o + a * ((b + 31) // 32)].reshape(int(a), int((b + 31) // 32)), int(b))
                          if a > 0 and b > 0 else np.zeros((int(a), int(b)), np.uint8) for o, a, b in zip(coff, Mi, Nj)]
        return out

    # ------------------------------------------------------------------ scoring
    def serra09_pairs(self, pairs, m=9, kappa=0.095, flags=0, want_crp=False, want_sw=False):
        """pairs: (P, 2) int array.  Returns dict with qmax_raw, dmax_raw (float32), oti (int32),
        shape (P, 2) = (M, N); with want_crp also 'crp' (list of uint8 (M, N) arrays in the
        reference's orientation) -- a debug path."""
        pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
        P = pairs.shape[0]
        if want_sw:
            flags |= FLAG_WANT_SW
        q = np.zeros(P, np.float32)
        d = np.zeros(P, np.float32)
        sw = np.zeros(P, np.float32)
        oti = np.zeros(P, np.int32)
        shape = np.zeros((P, 2), np.int32)
        crp_buf = crp_off = swapped = None
        if want_crp:
            L = self.lengths
            Mi = np.maximum(L[pairs[:, 0]].astype(np.int64) - m + 1, 0)
            Nj = np.maximum(L[pairs[:, 1]].astype(np.int64) - m + 1, 0)
            # room for either orientation
            words = np.maximum(Mi * ((Nj + 31) // 32), Nj * ((Mi + 31) // 32))
            crp_off = np.zeros(P, np.int64)
            if P > 1:
                crp_off[1:] = np.cumsum(words[:-1])
            crp_buf = np.zeros(max(int(words.sum()), 1), np.uint32)
            swapped = np.zeros(P, np.int32)
        rc = self.lib.acoss_serra09_pairs(self._h, _ptr(pairs, _c_i32p), P, int(m), float(kappa), int(flags),
                                          _ptr(q, _c_f32p), _ptr(d, _c_f32p), _ptr(sw, _c_f32p), _ptr(oti, _c_i32p),
                                          _ptr(shape, _c_i32p), _ptr(crp_buf, _c_u32p), _ptr(crp_off, _c_i64p),
                                          _ptr(swapped, _c_i32p))
        self._check(rc, "acoss_serra09_pairs")
        out = {"qmax_raw": q, "dmax_raw": d, "oti": oti, "shape": shape}
        if want_sw:
            out["sw_raw"] = sw
        if want_crp:
            crps = []
            for p in range(P):
                M, N = int(shape[p, 0]), int(shape[p, 1])
                if M <= 0 or N <= 0:
                    crps.append(np.zeros((max(M, 0), max(N, 0)), np.uint8))
                    continue
                r, c = (N, M) if swapped[p] else (M, N)
                wp = (c + 31) // 32
                B = unpack_bits(crp_buf[crp_off[p]:crp_off[p] + r * wp].reshape(r, wp), c)
                crps.append(np.ascontiguousarray(B.T) if swapped[p] else B)
            out["crp"] = crps
            out["swapped"] = swapped
        return out

    def align_dense(self, mats, kappa=0.095, flags=0, want_sw=False, want_crp=False):
        """mats: list of 2-D float32/float64 dissimilarity matrices (EarlySNF.py:86).  Returns
        (B, 3) float32 [qmax, dmax, sw] (+ list of uint8 CRPs with want_crp)."""
        B = len(mats)
        dt = np.float64 if any(a.dtype == np.float64 for a in mats) else np.float32
        rows = np.array([a.shape[0] for a in mats], np.int32)
        cols = np.array([a.shape[1] for a in mats], np.int32)
        sizes = rows.astype(np.int64) * cols
        offs = np.zeros(B, np.int64)
        if B > 1:
            offs[1:] = np.cumsum(sizes[:-1])
        flat = np.empty(max(int(sizes.sum()), 1), dt)
        for a, o, n in zip(mats, offs, sizes):
            flat[o:o + n] = np.asarray(a, dtype=dt).reshape(-1)
        if want_sw:
            flags |= FLAG_WANT_SW
        out = np.zeros((B, 3), np.float32)
        crp_buf = crp_off = None
        if want_crp:
            words = rows.astype(np.int64) * ((cols + 31) // 32)
            crp_off = np.zeros(B, np.int64)
            if B > 1:
                crp_off[1:] = np.cumsum(words[:-1])
            crp_buf = np.zeros(max(int(words.sum()), 1), np.uint32)
        rc = self.lib.acoss_align_dense(self._h, flat.ctypes.data_as(ctypes.c_void_p), F64 if dt == np.float64 else F32,
                                        _ptr(offs, _c_i64p), _ptr(rows, _c_i32p), _ptr(cols, _c_i32p), B,
                                        float(kappa), int(flags), _ptr(out, _c_f32p), _ptr(crp_buf, _c_u32p),
                                        _ptr(crp_off, _c_i64p))
        self._check(rc, "acoss_align_dense")
        if want_crp:
            crps = []
            for b in range(B):
                wp = (int(cols[b]) + 31) // 32
                crps.append(unpack_bits(crp_buf[crp_off[b]:crp_off[b] + int(rows[b]) * wp].reshape(int(rows[b]), wp),
                                        int(cols[b])))
            return out, crps
        return out

    def aggregate_songs(self, arrays, fac=40, aggregate="median", layout="frames_x_bins", zero_nonfinite=False,
                        want_gchroma=False):
        """Block aggregation of full-rate features on the device (Serra09.load_features,
        Serra09.py:101-110).  arrays: list of 2-D float32 or float64 arrays, all (n, bins)
        ("frames_x_bins": hpcp / crema) or all (bins, n) ("bins_x_frames": mfcc_htk).  Returns a
        list of (bins, ceil(n / fac)) arrays like librosa.util.sync(a.T or a, arange(0, n, fac))
        and, with want_gchroma, the (n_songs, bins) global chroma (Serra09.py:24-28)."""
        # groups of at most 256 MB of input: the page-locked staging buffers stay small and are allocated once
        budget, groups, cur, size = 256 << 20, [], [], 0
        for a in arrays:
            nb = int(np.asarray(a).size) * 8
            if cur and size + nb > budget:
                groups.append(cur)
                cur, size = [], 0
            cur.append(a)
            size += nb
        if len(groups) >= 1:
            groups.append(cur)
            res, gs, ms, nbytes = [], [], 0.0, 0.0
            for grp in groups:
                r = self.aggregate_songs(grp, fac, aggregate, layout, zero_nonfinite, want_gchroma)
                res += r[0] if want_gchroma else r
                if want_gchroma:
                    gs.append(r[1])
                a, b = self.last_aggregate_stats()
                ms, nbytes = ms + a, nbytes + b
            self._agg_stats = (ms, nbytes)
            if want_gchroma:
                dt = np.float64 if any(g.dtype == np.float64 for g in gs) else np.float32
                return res, np.concatenate([g.astype(dt) for g in gs], axis=0)
            return res
        S = len(arrays)
        dt = np.float64 if any(np.asarray(a).dtype == np.float64 for a in arrays) else np.float32
        lay = 0 if layout == "frames_x_bins" else 1
        bins = int(arrays[0].shape[1 if lay == 0 else 0]) if S else 0
        nfr = np.array([a.shape[0 if lay == 0 else 1] for a in arrays], np.int32)
        offs = np.zeros(S, np.int64)
        if S > 1:
            offs[1:] = np.cumsum(nfr[:-1].astype(np.int64))
        # staging buffers in page-locked memory, kept by the context and grown as needed (allocating them costs more
        # than the copy): the library then moves a chunk with one asynchronous copy at PCIe speed
        flat = self._staging("in", max(int(nfr.astype(np.int64).sum()) * max(bins, 1), 1), dt)
        for a, o, n in zip(arrays, offs, nfr):
            a = np.ascontiguousarray(a, dtype=dt)
            if a.shape[1 if lay == 0 else 0] != bins:
                raise ValueError("all songs must have the same number of bins")
            flat[o * bins:(o + int(n)) * bins] = a.reshape(-1)
        L = (nfr.astype(np.int64) + fac - 1) // fac
        ooffs = np.zeros(S, np.int64)
        if S > 1:
            ooffs[1:] = np.cumsum(L[:-1])
        out = self._staging("out", max(int(L.sum()) * max(bins, 1), 1), dt)
        g = np.empty((S, bins), dt) if want_gchroma else None
        rc = self.lib.acoss_aggregate_songs(self._h, flat.ctypes.data_as(ctypes.c_void_p), F64 if dt == np.float64 else F32,
                                            lay, _ptr(offs, _c_i64p), _ptr(nfr, _c_i32p), S, bins, int(fac),
                                            0 if aggregate == "median" else 1, 1 if zero_nonfinite else 0,
                                            out.ctypes.data_as(ctypes.c_void_p), _ptr(ooffs, _c_i64p),
                                            None if g is None else g.ctypes.data_as(ctypes.c_void_p))
        self._check(rc, "acoss_aggregate_songs")
        res = [np.array(out[o * bins:(o + int(l)) * bins].reshape(int(l), bins).T) for o, l in zip(ooffs, L)]   # copies: `out` is reused
        return (res, g) if want_gchroma else res

    def last_aggregate_stats(self):
        """(device ms of the aggregation kernel launches, bytes they moved) of the last aggregate_songs call."""
        if getattr(self, "_agg_stats", None) is not None:
            st, self._agg_stats = self._agg_stats, None
            return st
        out = np.zeros(2, np.float64)
        self._check(self.lib.acoss_last_aggregate_stats(self._h, out.ctypes.data_as(ctypes.POINTER(ctypes.c_double))),
                    "acoss_last_aggregate_stats")
        return float(out[0]), float(out[1])

    def _staging(self, which, n, dtype):
        """First n elements (as dtype) of a page-locked byte buffer owned by this context."""
        need = int(n) * np.dtype(dtype).itemsize
        bufs = self.__dict__.setdefault("_stage_bufs", {})
        buf = bufs.get(which)
        if buf is None or buf.nbytes < need:
            buf = _host_array((max(need + need // 4, 1 << 20),), np.uint8, pinned=True)
            bufs[which] = buf
        return buf[:need].view(dtype)

    def eval_ranks(self, D, cliques, device_ptr=None):
        """The rank loop of CoverAlgorithm.getEvalStatistics (CoverAlgorithm.py:345-382) on the device.
        D: (n, n) float32 score matrix (host), or None with device_ptr = address of such a matrix
        in device memory (a torch tensor's data_ptr()) and n given by the cliques.  cliques: list
        of lists of song indices, in the order of CoverAlgorithm.cliques.  Returns (iranks, Ks,
        n_rows): iranks = list, one int32 array of K - 1 ascending ranks per song of every clique
        with K >= 2 (clique-contiguous order, cliques by descending size), Ks = sorted sizes."""
        order, Ks, cstart, csize, offs, rows, total = eval_layout(cliques)
        n = int(order.size)
        out = np.zeros(max(total, 1), np.int32)
        if device_ptr is None:
            D = np.ascontiguousarray(D, dtype=np.float32)
            if D.shape != (n, n):
                raise ValueError("D must be (%d, %d)" % (n, n))
            ptr, flags = D.ctypes.data_as(ctypes.c_void_p), 0
        else:
            ptr, flags = ctypes.c_void_p(int(device_ptr)), FLAG_DEVICE_INPUT
        rc = self.lib.acoss_eval_ranks(self._h, ptr, n, _ptr(order, _c_i32p), _ptr(cstart, _c_i32p),
                                       _ptr(csize, _c_i32p), rows, _ptr(offs, _c_i64p), _ptr(out, _c_i32p), flags)
        self._check(rc, "acoss_eval_ranks")
        return [out[o:o + k - 1] for o, k in zip(offs[:rows], csize[:rows])], Ks, rows

    def last_eval_ms(self):
        out = np.zeros(1, np.float32)
        self._check(self.lib.acoss_last_eval_ms(self._h, _ptr(out, _c_f32p)), "acoss_last_eval_ms")
        return float(out[0])

    def dp_packed(self, plots, flags=0, want_sw=False):
        """plots: list of uint8 (rows, cols) binary matrices.  Returns (B, 3) [qmax, dmax, sw]."""
        B = len(plots)
        rows = np.array([a.shape[0] for a in plots], np.int32)
        cols = np.array([a.shape[1] for a in plots], np.int32)
        packed = [pack_bits(np.asarray(a, dtype=np.uint8)) for a in plots]
        words = np.array([p.size for p in packed], np.int64)
        offs = np.zeros(B, np.int64)
        if B > 1:
            offs[1:] = np.cumsum(words[:-1])
        buf = np.zeros(max(int(words.sum()), 1), np.uint32)
        for p, o in zip(packed, offs):
            buf[o:o + p.size] = p.reshape(-1)
        if want_sw:
            flags |= FLAG_WANT_SW
        out = np.zeros((B, 3), np.float32)
        rc = self.lib.acoss_dp_packed(self._h, _ptr(buf, _c_u32p), _ptr(offs, _c_i64p), _ptr(rows, _c_i32p),
                                      _ptr(cols, _c_i32p), B, int(flags), _ptr(out, _c_f32p))
        self._check(rc, "acoss_dp_packed")
        return out

    # pySeqAlign-compatible scalar calls (pySeqAlign.pyx:7-23)
    def _scalar(self, fn, S, a, b):
        S = np.ascontiguousarray(S, dtype=np.uint8).reshape(-1)
        if S.size < a * b:
            raise ValueError("S has %d elements, need %d" % (S.size, a * b))
        r = ctypes.c_float(0.0)
        self._check(fn(self._h, _ptr(S, _c_u8p), int(a), int(b), ctypes.byref(r)), fn.__name__)
        return float(r.value)

    def qmax(self, S, M, N):
        return self._scalar(self.lib.acoss_qmax, S, M, N)

    def dmax(self, S, M, N):
        return self._scalar(self.lib.acoss_dmax, S, M, N)

    def swconstrained(self, S, N, M):
        return self._scalar(self.lib.acoss_swconstrained, S, N, M)

    def qmax_dmax(self, S, M, N):
        S = np.ascontiguousarray(S, dtype=np.uint8).reshape(-1)
        q, d = ctypes.c_float(0.0), ctypes.c_float(0.0)
        self._check(self.lib.acoss_qmax_dmax(self._h, _ptr(S, _c_u8p), int(M), int(N), ctypes.byref(q), ctypes.byref(d)),
                    "acoss_qmax_dmax")
        return float(q.value), float(d.value)

    def set_option(self, name, value):
        self._check(self.lib.acoss_set_option(self._h, name.encode(), int(value)), "acoss_set_option")

    # ------------------------------------------------------------------ accounting
    def kernel_launches(self):
        return int(self.lib.acoss_kernel_launches(self._h))

    def reset_counters(self):
        self.lib.acoss_reset_counters(self._h)

    def bind_score_matrices(self, q_ptr, d_ptr, n, song_of_slot=None):
        """acoss_bind_score_matrices: q_ptr / d_ptr are DEVICE addresses of n x n float32 matrices (torch tensors'
        data_ptr(); 0 / None = not wanted, both = unbind); song_of_slot maps uploaded slots to matrix indices."""
        m = None if song_of_slot is None else np.ascontiguousarray(song_of_slot, dtype=np.int32)
        rc = self.lib.acoss_bind_score_matrices(self._h, ctypes.c_void_p(int(q_ptr)) if q_ptr else None,
                                                ctypes.c_void_p(int(d_ptr)) if d_ptr else None, int(n),
                                                _ptr(m, _c_i32p), 0 if m is None else int(m.size))
        self._check(rc, "acoss_bind_score_matrices")

    def debug_host_ms(self):
        out = np.zeros(6, np.float32)
        self._check(self.lib.acoss_debug_host_ms(self._h, _ptr(out, _c_f32p)), "acoss_debug_host_ms")
        return dict(zip(["call", "first_prep", "issue", "overlapped_prep", "wait", "results"], out.tolist()))

    def last_stage_ms(self):
        out = np.zeros(5, np.float32)
        self._check(self.lib.acoss_last_stage_ms(self._h, _ptr(out, _c_f32p)), "acoss_last_stage_ms")
        return dict(zip(["oti", "thresholds", "crp", "dp", "total"], out.tolist()))


def dp_packed_static(lib, device, crp_ptr, crp_offsets, rows, cols, out_ptr, flags=0, stream=0):
    """acoss_dp_packed_static: the DPs over packed plots that already live in DEVICE memory (crp_ptr / out_ptr are
    device addresses, e.g. torch tensors' data_ptr(); out holds 3 * B floats: qmax | dmax | sw), no allocation and no
    synchronisation inside the call.  Raises AcossError with the library's message on failure."""
    crp_offsets = np.ascontiguousarray(crp_offsets, dtype=np.int64)
    rows = np.ascontiguousarray(rows, dtype=np.int32)
    cols = np.ascontiguousarray(cols, dtype=np.int32)
    rc = lib.acoss_dp_packed_static(int(device), ctypes.c_void_p(int(crp_ptr)), _ptr(crp_offsets, _c_i64p), _ptr(rows, _c_i32p),
                                    _ptr(cols, _c_i32p), int(len(rows)), int(flags), ctypes.c_void_p(int(out_ptr)),
                                    ctypes.c_void_p(int(stream)) if stream else None)
    if rc != 0:
        raise AcossError("acoss_dp_packed_static failed (%d): %s" % (rc, lib.acoss_last_error().decode()))


def eval_layout(cliques):
    """Host bookkeeping of getEvalStatistics step 1 (CoverAlgorithm.py:345-357): cliques sorted by
    size (descending; np.argsort(-Ks) as the reference calls it), songs unrolled in that order.
    Returns order (position -> song), Ks, and for the leading positions that belong to cliques of
    >= 2 songs: clique start, clique size, output offsets; their count; total number of ranks."""
    cl = [list(c) for c in cliques]
    Ks = np.array([len(c) for c in cl], dtype=np.int64)
    idx = np.argsort(-Ks)
    Ks = Ks[idx]
    order = np.array([s for i in idx for s in cl[i]], dtype=np.int32)
    big = Ks >= 2
    # the reference stops at the first clique with fewer than two songs (:374-377)
    nbig = int(np.argmin(big)) if not big.all() else len(Ks)
    Kb = Ks[:nbig]
    starts = np.concatenate(([0], np.cumsum(Kb)))[:-1] if nbig else np.zeros(0, np.int64)
    cstart = np.repeat(starts, Kb).astype(np.int32)
    csize = np.repeat(Kb, Kb).astype(np.int32)
    rows = int(cstart.size)
    offs = np.zeros(max(rows, 1), np.int64)
    if rows > 1:
        offs[1:rows] = np.cumsum(csize[:-1].astype(np.int64) - 1)
    total = int((csize.astype(np.int64) - 1).sum())
    if rows == 0:
        cstart = np.zeros(1, np.int32)
        csize = np.zeros(1, np.int32)
    return order, Ks, cstart, csize, offs, rows, total


def pack_bits(B):
    """(rows, cols) 0/1 -> (rows, ceil(cols/32)) uint32, bit t of word w = B[:, 32w+t]."""
    B = np.asarray(B, dtype=np.uint8)
    rows, cols = B.shape
    wp = (cols + 31) // 32
    pad = np.zeros((rows, wp * 32), dtype=np.uint8)
    pad[:, :cols] = B != 0
    by = np.packbits(pad.reshape(rows, wp, 4, 8), axis=-1, bitorder="little").reshape(rows, wp, 4)
    return np.ascontiguousarray(by).view(np.uint32).reshape(rows, wp) if by.size else np.zeros((rows, wp), np.uint32)


def unpack_bits(P, cols):
    P = np.ascontiguousarray(P, dtype=np.uint32)
    rows, wp = P.shape
    by = P.view(np.uint8).reshape(rows, wp * 4)
    bits = np.unpackbits(by, axis=1, bitorder="little")
    return np.ascontiguousarray(bits[:, :cols])
