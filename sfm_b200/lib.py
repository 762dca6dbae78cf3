"""ctypes binding of libsfm_match.so (C ABI in include/sfm_match.h).

PyTorch is used only to own device memory and streams; every compute call goes
through the C ABI into hand-written sm_100a kernels.  There is no CPU or eager
fallback: if the library or a CUDA device is missing, calls raise.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

PKG = Path(__file__).resolve().parent
LIB_PATH = PKG / "libsfm_match.so"
ABI_VERSION = 1

QBLOCK = 256
TTILE = 128
INT32_MAX = 2**31 - 1

PAIR_OK, PAIR_TOO_FEW_FEATURES, PAIR_TOO_FEW_MATCHES, PAIR_NO_MODEL, PAIR_QUALITY_FAIL = range(5)
STATUS_NAMES = ("OK", "TOO_FEW_FEATURES", "TOO_FEW_MATCHES", "NO_MODEL", "QUALITY_FAIL")


class VerifyParams(C.Structure):
    """sfm_verify_params; defaults = the constants hard-coded in the reference
    (utils/find_matches.py:157, :185, :203, :274, :282)."""
    _fields_ = [
        ("ransac_thresh", C.c_double), ("confidence", C.c_double),
        ("max_iters", C.c_int32), ("min_matches", C.c_int32), ("seed", C.c_uint64),
        ("verify_thresh", C.c_double), ("min_inliers", C.c_int32), ("lo_iters", C.c_int32),
        ("min_ratio", C.c_double), ("max_error", C.c_double), ("min_spread", C.c_double),
        ("score_mode", C.c_int32), ("min_work", C.c_int32),
    ]


METRICS_DTYPE = np.dtype([
    ("total_matches", np.int32), ("inliers", np.int32), ("inlier_ratio", np.float32),
    ("reprojection_error", np.float32), ("symmetric_error", np.float32),
    ("well_distributed", np.int32), ("ransac_inliers", np.int32), ("ransac_iters", np.int32),
])

# every symbol include/sfm_match.h declares
EXPORTS = (
    "sfm_create", "sfm_destroy", "sfm_last_error", "sfm_default_params", "sfm_abi_version",
    "sfm_set_images", "sfm_knn2", "sfm_filter_matches", "sfm_verify_pairs", "sfm_verify_given_F",
    "sfm_ransac_points", "sfm_pack_results", "sfm_radius_match", "sfm_launch_count", "sfm_probe_int8_peak", "sfm_debug_dot_tile",
    "sfm_mutual_filter", "sfm_filter_matches_static", "sfm_debug_wide_tile", "sfm_debug_ts_tile", "sfm_debug_pair_tile", "sfm_debug_pair_counters", "sfm_debug_ts_rate_probe", "sfm_debug_handover_probe", "sfm_debug_set_knn_mode", "sfm_debug_fold_flag", "sfm_debug_tmem_ld_probe", "sfm_debug_set_knn_waits", "sfm_probe_int8_rate", "sfm_probe_fp64_rate", "sfm_debug_mma_handshake_probe", "sfm_debug_brute_rows", "sfm_debug_set_knn_probe_flags",
)

_lib = None


class SfmError(RuntimeError):
    pass


def load_library(path: Path | None = None) -> C.CDLL:
    """dlopen the in-tree library and declare the prototypes.  Raises if absent."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = Path(path) if path else LIB_PATH
    if not p.exists():
        raise SfmError(f"{p} is missing: run `python -m sfm_b200.build` (nvcc, sm_100a). "
                       "There is no CPU fallback for the matching path.")
    lib = C.CDLL(str(p))
    vp, i32, i64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_double
    lib.sfm_abi_version.restype = C.c_int
    lib.sfm_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.sfm_destroy.argtypes = [vp]
    lib.sfm_destroy.restype = None
    lib.sfm_last_error.argtypes = [vp]
    lib.sfm_last_error.restype = C.c_char_p
    lib.sfm_default_params.argtypes = [C.POINTER(VerifyParams)]
    lib.sfm_default_params.restype = None
    lib.sfm_set_images.argtypes = [vp, vp, vp, vp, i32, vp]
    lib.sfm_knn2.argtypes = [vp, vp, i32, vp, vp, vp]
    lib.sfm_filter_matches.argtypes = [vp, vp, i32, vp, vp, vp, vp, dbl, vp, vp, vp, vp, vp]
    lib.sfm_filter_matches_static.argtypes = [vp, vp, i32, vp, vp, dbl, vp, vp, vp, vp, vp]
    lib.sfm_mutual_filter.argtypes = [vp, vp, i32, vp, vp, vp, vp, vp, vp, vp]
    lib.sfm_verify_pairs.argtypes = [vp, vp, i32, vp, vp, vp, vp, C.POINTER(VerifyParams),
                                     vp, vp, vp, vp, vp, vp, vp, vp]
    lib.sfm_verify_given_F.argtypes = [vp, i32, vp, vp, vp, vp, C.POINTER(VerifyParams),
                                       vp, vp, vp, vp, vp]
    lib.sfm_ransac_points.argtypes = [vp, i32, vp, vp, vp, C.POINTER(VerifyParams),
                                      vp, vp, vp, vp, vp, vp]
    lib.sfm_pack_results.argtypes = [vp, i32, vp, vp, vp, i32] + [vp] * 15 + [vp]
    lib.sfm_radius_match.argtypes = [vp, i32, vp, vp, vp, vp, dbl, vp, vp, vp, i64, vp]
    lib.sfm_launch_count.argtypes = [vp]
    lib.sfm_launch_count.restype = i64
    lib.sfm_probe_int8_peak.argtypes = [vp, i32, C.POINTER(C.c_float), C.POINTER(C.c_double), vp]
    lib.sfm_debug_dot_tile.argtypes = [vp, i64, i64, vp, vp]
    lib.sfm_probe_int8_rate.argtypes = [vp, i32, i32, C.POINTER(C.c_float), C.POINTER(C.c_double), vp]
    lib.sfm_probe_fp64_rate.argtypes = [vp, i32, C.POINTER(C.c_float), C.POINTER(C.c_double), vp]
    lib.sfm_debug_wide_tile.argtypes = [vp, i64, i64, vp, vp]
    lib.sfm_debug_ts_tile.argtypes = [vp, i64, i64, vp, vp]
    lib.sfm_debug_pair_tile.argtypes = [vp, i64, i64, vp, vp]
    lib.sfm_debug_pair_counters.argtypes = [vp, C.POINTER(i64), vp]
    lib.sfm_debug_ts_rate_probe.argtypes = [vp, i32, i32, i32, i32, C.POINTER(C.c_float), C.POINTER(C.c_double), vp]
    lib.sfm_debug_set_knn_mode.argtypes = [vp, i32]
    lib.sfm_debug_handover_probe.argtypes = [vp, i32, i32, i32, i32, C.POINTER(i64), vp]
    lib.sfm_debug_set_knn_waits.argtypes = [vp, i32, i32]
    lib.sfm_debug_mma_handshake_probe.argtypes = [vp, i32, i32, i32, i32, C.POINTER(C.c_float), vp]
    lib.sfm_debug_set_knn_probe_flags.argtypes = [vp, i32]
    lib.sfm_debug_tmem_ld_probe.argtypes = [vp, i32, i32, i32, C.POINTER(i64), vp]
    lib.sfm_debug_fold_flag.argtypes = [vp, C.POINTER(i32), vp]
    lib.sfm_debug_brute_rows.argtypes = [vp, C.POINTER(i32), vp]
    for name in EXPORTS:
        getattr(lib, name)
    if lib.sfm_abi_version() != ABI_VERSION:
        raise SfmError("libsfm_match.so ABI version mismatch; rebuild")
    if path is None:
        _lib = lib
    return lib


def default_params(**overrides) -> VerifyParams:
    p = VerifyParams()
    load_library().sfm_default_params(C.byref(p))
    for k, v in overrides.items():
        if not hasattr(p, k):
            raise AttributeError(k)
        setattr(p, k, v)
    return p


def _h(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def _i64(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, dtype=np.int64))


def _pairs(a) -> np.ndarray:
    a = np.ascontiguousarray(np.asarray(a, dtype=np.int32)).reshape(-1, 2)
    return a


class Context:
    """One sfm_ctx: one GPU, one host thread."""

    def __init__(self, device: int = 0):
        import torch
        if not torch.cuda.is_available():
            raise SfmError("no CUDA device: the matching path has no CPU fallback")
        self.torch = torch
        self.lib = load_library()
        self.device = torch.device("cuda", device)
        h = C.c_void_p()
        if self.lib.sfm_create(device, C.byref(h)) != 0:
            raise SfmError(self.lib.sfm_last_error(None).decode())
        self.h = h
        self._keep = None
        self.offsets = None

    def close(self):
        if getattr(self, "h", None):
            self.lib.sfm_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- helpers
    def _check(self, rc: int):
        if rc != 0:
            raise SfmError(self.lib.sfm_last_error(self.h).decode())

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream)

    def _dev(self, t, dtype):
        torch = self.torch
        if t.device != self.device or t.dtype != dtype or not t.is_contiguous():
            raise SfmError(f"expected contiguous {dtype} tensor on {self.device}")
        if t.numel() == 0:
            # torch reports a NULL data pointer for an empty tensor; the C ABI takes NULL for "argument
            # missing", so a legitimately empty array (images without features) is given real storage
            if getattr(self, "_nil", None) is None:
                self._nil = torch.zeros(64, dtype=torch.uint8, device=self.device)
            return C.c_void_p(self._nil.data_ptr())
        return C.c_void_p(t.data_ptr())

    def empty(self, shape, dtype):
        shape = tuple(int(x) for x in (shape if isinstance(shape, (tuple, list)) else (shape,)))
        return self.torch.empty(shape, dtype=dtype, device=self.device)

    @property
    def launches(self) -> int:
        return int(self.lib.sfm_launch_count(self.h))

    # -- image set
    def set_images(self, desc, kps, offsets):
        """desc u8 [R,128] cuda, kps f32 [R,2] cuda or None, offsets int64 [N+1] (host)."""
        torch = self.torch
        off = _i64(offsets.cpu().numpy() if hasattr(offsets, "cpu") else offsets)
        if desc.dim() != 2 or desc.shape[1] != 128:
            raise SfmError("desc must be [rows, 128] uint8")
        if int(off[-1]) != desc.shape[0]:
            raise SfmError("offsets[-1] must equal the number of descriptor rows")
        if kps is not None and tuple(kps.shape) != (desc.shape[0], 2):
            raise SfmError("kps must be [rows, 2] float32")
        dptr = self._dev(desc, torch.uint8)              # (an empty set still passes real storage)
        kptr = self._dev(kps, torch.float32) if kps is not None else C.c_void_p(0)
        self._check(self.lib.sfm_set_images(self.h, dptr, kptr, _h(off), len(off) - 1, self._stream()))
        self._keep = (desc, kps)
        self.offsets = off
        self.counts = np.diff(off)

    def query_rows(self, pairs) -> np.ndarray:
        """int64 [P+1] output row offsets: pair p owns n_query(p) rows."""
        pairs = _pairs(pairs)
        ro = np.zeros(len(pairs) + 1, np.int64)
        np.cumsum(self.counts[pairs[:, 0]], out=ro[1:])
        return ro

    # -- K1
    def knn2(self, pairs, row_off=None, out=None):
        """Directed pairs -> i32 [rows,4] (idx1, idx2, d1^2, d2^2) on the device."""
        torch = self.torch
        pairs = _pairs(pairs)
        row_off = self.query_rows(pairs) if row_off is None else _i64(row_off)
        if out is None:
            out = self.empty((int(row_off[-1]), 4), torch.int32)
        self._check(self.lib.sfm_knn2(self.h, _h(pairs), len(pairs), _h(row_off),
                                      self._dev(out, torch.int32), self._stream()))
        return out, row_off

    # -- K2, mutual check in its sparse form
    def mutual_filter(self, pairs, row_off, count_host, mq, mt, md2, count):
        """Keep only the matches whose query row is the nearest row of the query image to their train
        descriptor; lists compacted in place, `count` rewritten.  count_host: int32 host copy of `count`."""
        torch = self.torch
        pairs = _pairs(pairs)
        row_off = _i64(row_off)
        ch = np.ascontiguousarray(np.asarray(count_host, np.int32))
        self._check(self.lib.sfm_mutual_filter(
            self.h, _h(pairs), len(pairs), _h(row_off), _h(ch), self._dev(mq, torch.int32),
            self._dev(mt, torch.int32), self._dev(md2, torch.int32), self._dev(count, torch.int32), self._stream()))

    def fold_flag(self) -> bool:
        """True when a registered descriptor's norm is beyond the folded kNN kernels' range (one sync)."""
        f = C.c_int32(0)
        self._check(self.lib.sfm_debug_fold_flag(self.h, C.byref(f), self._stream()))
        return f.value != 0

    # -- K2
    def filter_matches(self, pairs, row_off, knn, ratio=0.75, row_off_rev=None, knn_rev=None, out=None):
        torch = self.torch
        pairs = _pairs(pairs)
        row_off = _i64(row_off)
        n = int(row_off[-1])
        if out is None:
            out = (self.empty((n,), torch.int32), self.empty((n,), torch.int32),
                   self.empty((n,), torch.int32), self.empty((len(pairs),), torch.int32))
        mq, mt, md2, cnt = out
        rev_off = _i64(row_off_rev) if knn_rev is not None else None
        self._check(self.lib.sfm_filter_matches(
            self.h, _h(pairs), len(pairs), _h(row_off), self._dev(knn, torch.int32),
            _h(rev_off) if rev_off is not None else C.c_void_p(0),
            self._dev(knn_rev, torch.int32) if knn_rev is not None else C.c_void_p(0),
            float(ratio), self._dev(mq, torch.int32), self._dev(mt, torch.int32),
            self._dev(md2, torch.int32), self._dev(cnt, torch.int32), self._stream()))
        return mq, mt, md2, cnt

    def filter_matches_static(self, pairs, row_off, knn, ratio=0.75, out=None):
        """sfm_filter_matches_static: the ratio filter with its metadata in the library's static device
        array (the C entry point allocates nothing; pass `out` to keep this wrapper from allocating too)."""
        torch = self.torch
        pairs = _pairs(pairs)
        row_off = _i64(row_off)
        n = int(row_off[-1]) if len(row_off) else 0
        if out is None:
            out = (self.empty((n,), torch.int32), self.empty((n,), torch.int32),
                   self.empty((n,), torch.int32), self.empty((len(pairs),), torch.int32))
        mq, mt, md2, cnt = out
        self._check(self.lib.sfm_filter_matches_static(
            self.h, _h(pairs), len(pairs), _h(row_off), self._dev(knn, torch.int32), float(ratio),
            self._dev(mq, torch.int32), self._dev(mt, torch.int32), self._dev(md2, torch.int32),
            self._dev(cnt, torch.int32), self._stream()))
        return mq, mt, md2, cnt

    # -- K3 + K4
    def _verify_outputs(self, n_pairs, n_rows, need_F=True):
        torch = self.torch
        return dict(
            F=self.empty((n_pairs, 9), torch.float64) if need_F else None,
            mask=self.empty((n_rows,), torch.uint8),
            sym=self.empty((n_rows,), torch.float32),
            metrics=self.empty((n_pairs, METRICS_DTYPE.itemsize), torch.uint8),
            status=self.empty((n_pairs,), torch.int32),
        )

    def verify_pairs(self, pairs, row_off, mq, mt, count, params=None):
        torch = self.torch
        pairs = _pairs(pairs)
        row_off = _i64(row_off)
        n = int(row_off[-1])
        o = self._verify_outputs(len(pairs), n)
        o["pts1"] = self.empty((n, 2), torch.float32)
        o["pts2"] = self.empty((n, 2), torch.float32)
        prm = params if params is not None else default_params()
        self._check(self.lib.sfm_verify_pairs(
            self.h, _h(pairs), len(pairs), _h(row_off), self._dev(mq, torch.int32),
            self._dev(mt, torch.int32), self._dev(count, torch.int32), C.byref(prm),
            self._dev(o["pts1"], torch.float32), self._dev(o["pts2"], torch.float32),
            self._dev(o["F"], torch.float64), self._dev(o["mask"], torch.uint8),
            self._dev(o["sym"], torch.float32), self._dev(o["metrics"], torch.uint8),
            self._dev(o["status"], torch.int32), self._stream()))
        return o

    def verify_given_F(self, off, pts1, pts2, F, params=None):
        torch = self.torch
        off = _i64(off)
        o = self._verify_outputs(len(off) - 1, int(off[-1]), need_F=False)
        prm = params if params is not None else default_params()
        self._check(self.lib.sfm_verify_given_F(
            self.h, len(off) - 1, _h(off), self._dev(pts1, torch.float32),
            self._dev(pts2, torch.float32), self._dev(F, torch.float64), C.byref(prm),
            self._dev(o["mask"], torch.uint8), self._dev(o["sym"], torch.float32),
            self._dev(o["metrics"], torch.uint8), self._dev(o["status"], torch.int32), self._stream()))
        return o

    def ransac_points(self, off, pts1, pts2, params=None):
        torch = self.torch
        off = _i64(off)
        o = self._verify_outputs(len(off) - 1, int(off[-1]))
        prm = params if params is not None else default_params()
        self._check(self.lib.sfm_ransac_points(
            self.h, len(off) - 1, _h(off), self._dev(pts1, torch.float32),
            self._dev(pts2, torch.float32), C.byref(prm), self._dev(o["F"], torch.float64),
            self._dev(o["mask"], torch.uint8), self._dev(o["sym"], torch.float32),
            self._dev(o["metrics"], torch.uint8), self._dev(o["status"], torch.int32), self._stream()))
        return o

    def pack_results(self, row_off, count, status, only_ok, **arrays):
        """Gather the valid entries of the kept pairs into dense device arrays.

        arrays: any of mq, mt, md2 (int32), mask (uint8), sym, pts1, pts2 (float32).
        Returns (pack_off int64 [P+1] device tensor, dict of dense tensors)."""
        torch = self.torch
        row_off = _i64(row_off)
        n_pairs = len(row_off) - 1
        dt = dict(mq=torch.int32, mt=torch.int32, md2=torch.int32, mask=torch.uint8,
                  sym=torch.float32, pts1=torch.float32, pts2=torch.float32)
        pack_off = self.empty((n_pairs + 1,), torch.int64)
        src, dst, outs = [], [], {}
        for name in ("mq", "mt", "md2", "mask", "sym", "pts1", "pts2"):
            t = arrays.get(name)
            if t is None:
                src.append(C.c_void_p(0))
                dst.append(C.c_void_p(0))
            else:
                o = torch.empty_like(t)
                outs[name] = o
                src.append(self._dev(t, dt[name]))
                dst.append(self._dev(o, dt[name]))
        self._check(self.lib.sfm_pack_results(
            self.h, n_pairs, _h(row_off), self._dev(count, torch.int32),
            self._dev(status, torch.int32) if status is not None else C.c_void_p(0),
            1 if only_ok else 0, *src, C.c_void_p(pack_off.data_ptr()), *dst, self._stream()))
        return pack_off, outs

    # -- 2-D radius matching (find_2d3d_matches)
    def radius_match(self, track_off, tracks, pt_off, pts, threshold):
        """tracks f64 [rows,2], pts f64 [n,2] on the device; track_off / pt_off int64 [S+1] (host).

        Returns (seg_off int64 [S+1] numpy, track_idx int32 cuda, pt_idx int32 cuda): for segment s the
        entries [seg_off[s], seg_off[s+1]) are np.where(dist < threshold) of its tracks x points."""
        torch = self.torch
        track_off, pt_off = _i64(track_off), _i64(pt_off)
        n_seg = len(track_off) - 1
        if len(pt_off) != n_seg + 1:
            raise SfmError("track_off and pt_off must have the same length")
        if tuple(tracks.shape) != (int(track_off[-1]), 2) or tuple(pts.shape) != (int(pt_off[-1]), 2):
            raise SfmError("tracks / pts must be [rows, 2] matching the offsets")
        tp = self._dev(tracks, torch.float64) if tracks.numel() else C.c_void_p(0)
        pp = self._dev(pts, torch.float64) if pts.numel() else C.c_void_p(0)
        seg_off = np.zeros(n_seg + 1, np.int64)
        args = (self.h, n_seg, _h(track_off), tp, _h(pt_off), pp, float(threshold), _h(seg_off))
        cap = max(4 * int(track_off[-1]), 1024)      # a track matches a handful of points at most
        while True:
            out_t = self.empty((cap,), torch.int32)
            out_p = self.empty((cap,), torch.int32)
            self._check(self.lib.sfm_radius_match(*args, C.c_void_p(out_t.data_ptr()), C.c_void_p(out_p.data_ptr()),
                                                  cap, self._stream()))
            total = int(seg_off[-1])
            if total <= cap:
                return seg_off, out_t[:total], out_p[:total]
            cap = total

    # -- measurement / debug
    def probe_int8_peak(self, iters=4096):
        ms, ops = C.c_float(), C.c_double()
        self._check(self.lib.sfm_probe_int8_peak(self.h, iters, C.byref(ms), C.byref(ops), self._stream()))
        return ops.value / (ms.value * 1e-3) / 1e12, ms.value   # TOP/s, ms

    def probe_int8_rate(self, tile_n, iters=4096):
        ms, ops = C.c_float(), C.c_double()
        self._check(self.lib.sfm_probe_int8_rate(self.h, iters, tile_n, C.byref(ms), C.byref(ops), self._stream()))
        return ops.value / (ms.value * 1e-3) / 1e12, ms.value

    def debug_dot_tile(self, q_row, t_row):
        out = self.empty((128, 128), self.torch.int32)
        self._check(self.lib.sfm_debug_dot_tile(self.h, q_row, t_row, self._dev(out, self.torch.int32),
                                                self._stream()))
        return out


    def debug_wide_tile(self, q_row, t_row):
        out = self.empty((128, 256), self.torch.int32)
        self._check(self.lib.sfm_debug_wide_tile(self.h, q_row, t_row, self._dev(out, self.torch.int32),
                                                 self._stream()))
        return out

    def debug_ts_tile(self, q_row, t_row):
        out = self.empty((128, 128), self.torch.int32)
        self._check(self.lib.sfm_debug_ts_tile(self.h, q_row, t_row, self._dev(out, self.torch.int32),
                                               self._stream()))
        return out

    def debug_pair_tile(self, q_row, t_row):
        out = self.empty((256, 128), self.torch.int32)
        self._check(self.lib.sfm_debug_pair_tile(self.h, q_row, t_row, self._dev(out, self.torch.int32),
                                                 self._stream()))
        return out

    def pair_counters(self):
        buf = (C.c_int64 * 8)()
        self._check(self.lib.sfm_debug_pair_counters(self.h, buf, self._stream()))
        return list(buf)

    def ts_rate_probe(self, tile_n, ts=1, commit=0, groups=8192):
        """-> TOP/s of groups of five 128 x tile_n x 32 u8 MMAs (A in TMEM when ts, else shared memory)."""
        ms, ops = C.c_float(), C.c_double()
        self._check(self.lib.sfm_debug_ts_rate_probe(self.h, groups, tile_n, int(ts), int(commit), C.byref(ms),
                                                     C.byref(ops), self._stream()))
        return ops.value / (ms.value * 1e-3) / 1e12

    def tmem_ld_probe(self, warps, variant, iters=4096):
        """-> bytes of TMEM read per cycle per SM for the given warp count / load shape."""
        cyc = C.c_int64(0)
        self._check(self.lib.sfm_debug_tmem_ld_probe(self.h, warps, variant, iters, C.byref(cyc), self._stream()))
        per = {0: 4096, 1: 4096, 2: 8192, 3: 4096, 4: 2048, 5: 2048, 6: 2048}[variant]
        return per * iters * warps / cyc.value, cyc.value

    def probe_fp64_rate(self, iters=4096):
        """-> f64 instructions per second of an all-SM DFMA loop (the radius matcher's roofline)."""
        ms, n = C.c_float(), C.c_double()
        self._check(self.lib.sfm_probe_fp64_rate(self.h, iters, C.byref(ms), C.byref(n), self._stream()))
        return n.value / (ms.value * 1e-3)

    def set_knn_mode(self, mode: int):
        """0 auto, 1 generic, 5 wide, 12 CTA pair, 15 wide + skip, 16 wide + lean tracker, 17 ping-pong;
        6 / 7 / 13 / 14 / 18 / 19 timing probes (sfm_match.h)."""
        self._check(self.lib.sfm_debug_set_knn_mode(self.h, mode))

    def set_knn_waits(self, wait_ns: int, epi_wait_ns: int):
        self._check(self.lib.sfm_debug_set_knn_waits(self.h, wait_ns, epi_wait_ns))

    def mma_handshake_probe(self, k_per_acc, flags, consumers, steps=4096):
        """-> TOP/s of the issue pattern (2 x k_per_acc 128x128x32 MMAs per step)."""
        ms = C.c_float()
        self._check(self.lib.sfm_debug_mma_handshake_probe(self.h, steps, k_per_acc, flags, consumers,
                                                           C.byref(ms), self._stream()))
        ops = 2.0 * 128 * 128 * 32 * 2 * k_per_acc * steps * 148
        return ops / (ms.value * 1e-3) / 1e12

    def brute_rows(self) -> int:
        f = C.c_int32(-1)
        self._check(self.lib.sfm_debug_brute_rows(self.h, C.byref(f), self._stream()))
        return f.value

    def set_knn_probe_flags(self, flags: int):
        self._check(self.lib.sfm_debug_set_knn_probe_flags(self.h, flags))

    def fold_flag(self) -> int:
        f = C.c_int32(-1)
        self._check(self.lib.sfm_debug_fold_flag(self.h, C.byref(f), self._stream()))
        return f.value


def metrics_to_numpy(t) -> np.ndarray:
    """uint8 [P, 32] device tensor (or its host copy as numpy) -> structured array (METRICS_DTYPE)."""
    a = t if isinstance(t, np.ndarray) else t.cpu().numpy()
    return a.view(METRICS_DTYPE).reshape(-1)
