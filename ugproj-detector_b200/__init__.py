"""ugproj-detector_b200 -- host-side mirror of the reference's detector interface over libugdet.so.

The product is the CUDA library (csrc/, C ABI in include/ugdet.h).  This module is the thin
Python face of it used by tests and bench.py; `FaceDetector` mirrors ugproj::FaceDetector
(reference src/tracker/detector.h:10-19): constructed from a loaded cascade + the [detection]
config section, `detectFaces(frame, rects)` APPENDS to the caller's list (detector.cc:41).

There is no CPU path here: importing works without a GPU (the library only needs libcudart's
static stub), but creating a detector without a CUDA device raises UgdError(UGD_ERR_CUDA).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _build

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

OK, ERR_ARG, ERR_IO, ERR_PARSE, ERR_UNSUPPORTED, ERR_CUDA, ERR_CAPACITY, ERR_NOMEM, ERR_TRUNCATED = range(9)
K_NAMES = ("k_preprocess", "k_equalize", "k_resize", "k_colsum", "k_carry", "k_integral", "k_tilted", "k_scan",
           "k_group", "k_area")


class Rect(C.Structure):
    _fields_ = [("x", C.c_int32), ("y", C.c_int32), ("width", C.c_int32), ("height", C.c_int32)]


class Params(C.Structure):
    _fields_ = [("scale_factor", C.c_double), ("min_neighbors", C.c_int32), ("min_w", C.c_int32),
                ("min_h", C.c_int32), ("max_w", C.c_int32), ("max_h", C.c_int32),
                ("skin_lower", C.c_uint8 * 3), ("skin_upper", C.c_uint8 * 3), ("_pad", C.c_uint8 * 2),
                ("dense_stages", C.c_int32), ("max_raw_per_frame", C.c_int32), ("legacy_24", C.c_int32)]


class CascadeInfo(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("win_w", "win_h", "nstages", "ntrees", "nnodes", "max_tree_nodes",
                                         "old_format", "sum_exact", "has_tilted")]


class Counters(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("frames", "levels", "level_pixels", "grid_windows", "windows_evaluated",
                                         "lowvar", "queued", "raw", "grouped", "near_threshold",
                                         "weak_evals")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


class RectLists:
    """The per-frame rect lists of one batch call, kept as the two arrays the C ABI filled -- rects [frames, cap, 4] and
    counts [frames] -- so that a call over hundreds of frames creates no per-frame Python objects until somebody looks.
    Behaves like a list of [n_i, 4] int32 arrays (len, indexing, iteration)."""
    __slots__ = ("rects", "counts", "start", "n")

    def __init__(self, rects, counts, start=0, n=None):
        self.rects, self.counts, self.start = rects, counts, start
        self.n = len(counts) - start if n is None else n

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(self.n))]
        if i < 0:
            i += self.n
        if not 0 <= i < self.n:
            raise IndexError(i)
        return self.rects[self.start + i, :min(int(self.counts[self.start + i]), self.rects.shape[1])]

    def __iter__(self):
        return (self[i] for i in range(self.n))

    def __add__(self, other):
        return list(self) + list(other)

    def total(self):
        return int(self.counts[self.start:self.start + self.n].sum())


class UgdError(RuntimeError):
    def __init__(self, code, detail=""):
        self.code = code
        msg = lib().ugd_strerror(code).decode()
        super().__init__(f"ugdet error {code} ({msg})" + (f": {detail}" if detail else ""))


def lib_path() -> str:
    return _build.LIB


def lib():
    """Loads libugdet.so (building it in-tree with nvcc if it is missing or stale)."""
    global _LIB
    if _LIB is None:
        path = os.environ.get("UGD_LIB") or _build.build()     # UGD_LIB: A/B a differently built library
        L = C.CDLL(path)
        L.ugd_strerror.restype = C.c_char_p
        L.ugd_last_error.restype = C.c_char_p
        L.ugd_version.restype = C.c_char_p
        L.ugd_kernel_name.restype = C.c_char_p
        L.ugd_cascade_free.restype = None
        L.ugd_destroy.restype = None
        L.ugd_params_default.restype = None
        L.ugd_last_error.argtypes = [C.c_void_p]
        L.ugd_destroy.argtypes = [C.c_void_p]
        L.ugd_cascade_free.argtypes = [C.c_void_p]
        L.ugd_shard_range.restype = None
        _LIB = L
    return _LIB


def default_params(**kw) -> Params:
    p = Params()
    lib().ugd_params_default(C.byref(p))
    for k, v in kw.items():
        if k in ("skin_lower", "skin_upper"):
            setattr(p, k, (C.c_uint8 * 3)(*v))
        elif k == "min_size":
            p.min_w, p.min_h = v
        elif k == "max_size":
            p.max_w, p.max_h = v
        else:
            setattr(p, k, v)
    return p


def load_config(cfg_path: str):
    """[detection] section of a reference cfg -> (Params, cascade_path, (scan_w, scan_h))."""
    p = Params()
    buf = C.create_string_buffer(4096)
    sw, sh = C.c_int32(), C.c_int32()
    rc = lib().ugd_config_load(cfg_path.encode(), C.byref(p), buf, C.c_size_t(4096), C.byref(sw), C.byref(sh))
    if rc:
        raise UgdError(rc, cfg_path)
    return p, buf.value.decode(), (sw.value, sh.value)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def comm_unique_id() -> bytes:
    """ncclGetUniqueId through the library (the root calls it; every rank passes the bytes to Detector.comm_init)."""
    buf = C.create_string_buffer(128)
    rc = lib().ugd_comm_unique_id(buf)
    if rc:
        raise UgdError(rc, "ugd_comm_unique_id")
    return buf.raw


def shard_range(nframes: int, rank: int, world: int):
    lo, hi = C.c_int(), C.c_int()
    lib().ugd_shard_range(nframes, rank, world, C.byref(lo), C.byref(hi))
    return lo.value, hi.value


class Cascade:
    """== cv::CascadeClassifier after load() (reference src/file_io.cc:11)."""

    def __init__(self, path: str):
        self._h = C.c_void_p()
        rc = lib().ugd_cascade_load_xml(path.encode(), C.byref(self._h))
        if rc:
            raise UgdError(rc, path)
        self.path = path
        self.info = CascadeInfo()
        lib().ugd_cascade_get_info(self._h, C.byref(self.info))

    def dump(self):
        i = self.info
        d = dict(stage_first_tree=np.zeros(i.nstages, np.int32), stage_ntrees=np.zeros(i.nstages, np.int32),
                 stage_thr=np.zeros(i.nstages, np.float32), tree_first_node=np.zeros(i.ntrees, np.int32),
                 tree_nnodes=np.zeros(i.ntrees, np.int32), node_rects=np.zeros((i.nnodes, 3, 4), np.int32),
                 node_weights=np.zeros((i.nnodes, 3), np.float32), node_thr=np.zeros(i.nnodes, np.float32),
                 node_left=np.zeros(i.nnodes, np.int32), node_right=np.zeros(i.nnodes, np.int32),
                 node_left_val=np.zeros(i.nnodes, np.float32), node_right_val=np.zeros(i.nnodes, np.float32),
                 node_tilted=np.zeros(i.nnodes, np.int32))
        rc = lib().ugd_cascade_dump(self._h, *[_p(v) for v in d.values()])
        if rc:
            raise UgdError(rc)
        return d

    def __del__(self):
        if getattr(self, "_h", None) and _LIB is not None:
            _LIB.ugd_cascade_free(self._h)
            self._h = None


def plan_query(cascade: "Cascade", params: Params | None, w: int, h: int):
    """Pyramid geometry (host arithmetic only, no GPU): scales, level sizes, steps, totals."""
    n = C.c_int(); gw = C.c_int64(); lp = C.c_int64()
    pp = C.byref(params) if params is not None else None
    rc = lib().ugd_plan_query(cascade._h, pp, w, h, C.byref(n), None, None, None, None, 0, C.byref(gw), C.byref(lp))
    if rc:
        raise UgdError(rc)
    sc = np.zeros(n.value, np.float32)
    lw = np.zeros(n.value, np.int32); lh = np.zeros(n.value, np.int32); st = np.zeros(n.value, np.int32)
    lib().ugd_plan_query(cascade._h, pp, w, h, C.byref(n), _p(sc), _p(lw), _p(lh), _p(st), n.value, None, None)
    return dict(scales=sc, lw=lw, lh=lh, step=st, grid_windows=gw.value, level_pixels=lp.value)


class Detector:
    """Owns one ugd_detector.  All frames are numpy uint8 arrays (or raw device pointers)."""

    def __init__(self, cascade: Cascade, params: Params | None = None, device: int = 0, max_size=(1280, 720),
                 max_batch: int = 1):
        self.cascade = cascade
        self.params = params if params is not None else default_params()
        self._h = C.c_void_p()
        rc = lib().ugd_create(cascade._h, C.byref(self.params), device, max_size[0], max_size[1], max_batch,
                              C.byref(self._h))
        if rc:
            raise UgdError(rc, "ugd_create")
        self.max_batch = max_batch

    def _ck(self, rc):
        if rc:
            raise UgdError(rc, lib().ugd_last_error(self._h).decode())

    def close(self):
        if getattr(self, "_h", None) and _LIB is not None:
            _LIB.ugd_destroy(self._h)
            self._h = None

    __del__ = close

    def set_stream(self, cuda_stream: int):
        self._ck(lib().ugd_set_stream(self._h, C.c_void_p(cuda_stream)))

    # -- single frame -----------------------------------------------------------------------------
    def detect_bgr(self, bgr: np.ndarray, cap: int = 1024) -> np.ndarray:
        assert bgr.dtype == np.uint8 and bgr.ndim == 3 and bgr.shape[2] == 3
        bgr = np.ascontiguousarray(bgr)
        h, w = bgr.shape[:2]
        return self._single(lib().ugd_detect_bgr, bgr, w, h, w * 3, cap)

    def _single(self, fn, frame, w, h, stride, cap):
        """One frame; a list longer than `cap` (UGD_ERR_TRUNCATED) is fetched again with the true count, so the
        result is never silently cut -- the reference's std::vector is unbounded (detector.cc:41)."""
        while True:
            out = (Rect * cap)()
            n = C.c_int()
            rc = fn(self._h, _p(frame), w, h, C.c_size_t(stride), out, cap, C.byref(n))
            if rc == ERR_TRUNCATED and n.value > cap:
                cap = n.value
                continue
            self._ck(rc)
            return _rects_np(out, min(n.value, cap))

    def detect_gray(self, gray: np.ndarray, cap: int = 1024) -> np.ndarray:
        assert gray.dtype == np.uint8 and gray.ndim == 2
        gray = np.ascontiguousarray(gray)
        h, w = gray.shape
        return self._single(lib().ugd_detect_gray, gray, w, h, w, cap)

    # -- batch ------------------------------------------------------------------------------------
    def detect_batch(self, frames: np.ndarray, cap: int = 256):
        """frames: [N,H,W] gray or [N,H,W,3] BGR host array -> list of [n_i,4] int32 arrays."""
        assert frames.dtype == np.uint8 and frames.ndim in (3, 4)
        frames = np.ascontiguousarray(frames)
        n, h, w = frames.shape[:3]
        return self._batch(_p(frames), 0, frames.ndim == 4, n, w, h, cap)

    def submit_batch(self, frames: np.ndarray, cap: int = 256) -> None:
        """Prefetching form (8(f) N2): enqueue a host batch and return at once; wait_batch() delivers the lists."""
        assert frames.dtype == np.uint8 and frames.ndim in (3, 4)
        self._pending = (np.ascontiguousarray(frames), cap)            # the frames must stay alive until wait_batch
        f = self._pending[0]
        n, h, w = f.shape[:3]
        self._ck(lib().ugd_submit_batch(self._h, _p(f), 0, int(f.ndim == 4), n, w, h, cap))

    def wait_batch(self):
        f, cap = self._pending
        n = f.shape[0]
        out = np.zeros((n, cap, 4), np.int32)
        cnt = np.zeros(n, np.int32)
        self._ck(lib().ugd_wait_batch(self._h, _p(out), _p(cnt)))
        self._pending = None
        return RectLists(out, cnt)

    def detect_batch_scaled(self, frames: np.ndarray, scan_size, cap: int = 256):
        """frames: [N,H,W,3] full-size BGR host array; scan_size = (w, h) as cfg scan.frame_size.  INTER_AREA
        decimation (face_tracker.cc:112-114) + detectFaces on the device; rects in scan-frame coordinates."""
        assert frames.dtype == np.uint8 and frames.ndim == 4 and frames.shape[3] == 3
        frames = np.ascontiguousarray(frames)
        n, h, w = frames.shape[:3]
        return self.detect_batch_scaled_ptr(_p(frames), False, n, w, h, scan_size, cap)

    def detect_batch_scaled_ptr(self, ptr, is_device: bool, n: int, w: int, h: int, scan_size, cap: int = 256):
        if not isinstance(ptr, C.c_void_p):
            ptr = C.c_void_p(ptr)
        out = np.empty((n, cap, 4), np.int32)
        cnt = np.zeros(n, np.int32)
        self._ck(lib().ugd_detect_batch_scaled(self._h, ptr, int(is_device), n, w, h, int(scan_size[0]), int(scan_size[1]),
                                               _p(out), cap, _p(cnt)))
        return RectLists(out, cnt)

    def crop_faces(self, dev_ptr: int, n: int, w: int, h: int, channels: int, lists: "RectLists"):
        """Face patches of device-resident frames (== the cv::Mat ROI copies of face_tracker.cc:296-302):
        list (per frame) of lists of [height, width(, 3)] uint8 arrays, in the order of `lists`."""
        cap = lists.rects.shape[1]
        rects = np.ascontiguousarray(lists.rects[lists.start:lists.start + n])
        counts = np.ascontiguousarray(lists.counts[lists.start:lists.start + n].astype(np.int32))
        offs = np.zeros(n * cap, np.uint64)
        total = C.c_size_t()
        rc = lib().ugd_crop_faces(self._h, C.c_void_p(dev_ptr), n, w, h, channels, _p(rects), _p(counts), cap, None, C.c_size_t(0),
                                  _p(offs), C.byref(total))
        if rc not in (OK, ERR_CAPACITY):
            self._ck(rc)
        buf = np.zeros(max(total.value, 1), np.uint8)
        self._ck(lib().ugd_crop_faces(self._h, C.c_void_p(dev_ptr), n, w, h, channels, _p(rects), _p(counts), cap, _p(buf),
                                      C.c_size_t(buf.size), _p(offs), C.byref(total)))
        out = []
        for f in range(n):
            row = []
            for i in range(min(int(counts[f]), cap)):
                x, y, rw, rh = (int(v) for v in rects[f, i])
                o = int(offs[f * cap + i])
                row.append(buf[o:o + rw * rh * channels].reshape((rh, rw, 3) if channels == 3 else (rh, rw)).copy())
            out.append(row)
        return out

    def resize_area(self, img: np.ndarray, dw: int, dh: int) -> np.ndarray:
        """cv2.resize(img, (dw, dh), interpolation=INTER_AREA) on the device (decimation only)."""
        img = np.ascontiguousarray(img, np.uint8)
        h, w = img.shape[:2]
        cn = 1 if img.ndim == 2 else img.shape[2]
        out = np.zeros((dh, dw) if img.ndim == 2 else (dh, dw, cn), np.uint8)
        self._ck(lib().ugd_resize_area(self._h, _p(img), w, h, cn, C.c_size_t(0), _p(out), dw, dh))
        return out

    def detect_batch_device(self, dev_ptr: int, is_bgr: bool, n: int, w: int, h: int, cap: int = 256):
        return self._batch(C.c_void_p(dev_ptr), 1, is_bgr, n, w, h, cap)

    def _batch(self, ptr, is_device, is_bgr, n, w, h, cap):
        while True:
            out = np.empty((max(n, 1), cap, 4), np.int32)            # fresh arrays: the result owns them, nothing is copied
            cnt = np.zeros(max(n, 1), np.int32)
            rc = lib().ugd_detect_batch(self._h, ptr, is_device, int(is_bgr), n, w, h, _p(out), cap, _p(cnt))
            if rc == ERR_TRUNCATED and int(cnt.max()) > cap:         # never cut a list silently: run again with room
                cap = int(cnt.max())
                continue
            self._ck(rc)
            return RectLists(out, cnt, 0, n)

    # -- multi-GPU: frames sharded over ranks, lists gathered to the root inside the library ----------
    def comm_init(self, comm_id: bytes, rank: int, world: int):
        """ncclCommInitRank on this detector's device; comm_id = comm_unique_id() of the root, handed to every rank."""
        assert len(comm_id) == 128
        self._ck(lib().ugd_comm_init(self._h, C.c_char_p(comm_id), rank, world))
        self.rank, self.world = rank, world           # cached: detect_batch_sharded is on the timed path

    def comm_info(self):
        r, w, v = C.c_int(), C.c_int(), C.c_int()
        self._ck(lib().ugd_comm_info(self._h, C.byref(r), C.byref(w), C.byref(v)))
        return r.value, w.value, v.value

    def detect_batch_sharded(self, ptr, is_device: bool, is_bgr: bool, n: int, w: int, h: int, slot_frames: int,
                             root: int = 0, cap: int = 256):
        """This rank's n frames (host array pointer or device pointer) + the NCCL gather to `root`.
        Root: list over ranks of lists over that rank's frames of [k,4] arrays; other ranks: their own lists."""
        if not isinstance(ptr, C.c_void_p):
            ptr = C.c_void_p(ptr)
        rank, world = getattr(self, "rank", 0), getattr(self, "world", 1)
        nslots = world * slot_frames if rank == root else n
        out = np.empty((max(nslots, 1), cap, 4), np.int32)           # fresh arrays: the result owns them, nothing is copied
        cnt = np.zeros(max(nslots, 1), np.int32)
        self._ck(lib().ugd_detect_batch_sharded(self._h, ptr, int(is_device), int(is_bgr), n, w, h, slot_frames, root,
                                                _p(out), cap, _p(cnt)))
        if rank != root:
            return RectLists(out, cnt, 0, n)
        # a rank's frames fill its slots from the front; the slots it did not use carry -1
        return [RectLists(out, cnt, r * slot_frames, int((cnt[r * slot_frames:(r + 1) * slot_frames] >= 0).sum())) for r in range(world)]

    # -- observability ----------------------------------------------------------------------------
    def counters(self) -> dict:
        c = Counters()
        self._ck(lib().ugd_get_counters(self._h, C.byref(c)))
        return c.as_dict()

    def set_profiling(self, on: bool):
        self._ck(lib().ugd_set_profiling(self._h, int(on)))

    def kernel_times(self):
        t = (C.c_float * len(K_NAMES))()
        l = (C.c_int64 * len(K_NAMES))()
        self._ck(lib().ugd_get_kernel_times(self._h, t, l))
        return {K_NAMES[i]: (float(t[i]), int(l[i])) for i in range(len(K_NAMES))}

    def plan_info(self, w, h):
        n = C.c_int()
        self._ck(lib().ugd_plan_info(self._h, w, h, C.byref(n), None, None, None, None, 0))
        sc = np.zeros(n.value, np.float32)
        lw = np.zeros(n.value, np.int32); lh = np.zeros(n.value, np.int32); st = np.zeros(n.value, np.int32)
        self._ck(lib().ugd_plan_info(self._h, w, h, C.byref(n), _p(sc), _p(lw), _p(lh), _p(st), n.value))
        return sc, lw, lh, st

    def raw(self, cap: int = 1 << 16):
        out = (Rect * cap)()
        sums = np.zeros(cap, np.float64)
        n = C.c_int()
        self._ck(lib().ugd_get_raw(self._h, out, _p(sums), cap, C.byref(n)))
        m = min(n.value, cap)
        return _rects_np(out, m), sums[:m].copy()

    def set_stage_map(self, on: bool):
        self._ck(lib().ugd_set_stage_map(self._h, int(on)))

    def stage_map(self):
        n = C.c_size_t()
        self._ck(lib().ugd_get_stage_map(self._h, None, C.c_size_t(0), C.byref(n)))
        m = np.zeros(n.value, np.int16)
        self._ck(lib().ugd_get_stage_map(self._h, _p(m), C.c_size_t(n.value), C.byref(n)))
        return m

    def preprocessed(self, w, h):
        g = np.zeros((h, w), np.uint8); m = np.zeros((h, w), np.uint8)
        self._ck(lib().ugd_get_preprocessed(self._h, _p(g), _p(m)))
        return g, m

    def level(self, idx, lw, lh):
        img = np.zeros((lh, lw), np.uint8)
        S = np.zeros((lh + 1, lw + 1), np.int32); SQ = np.zeros((lh + 1, lw + 1), np.int32)
        self._ck(lib().ugd_get_level(self._h, idx, _p(img), _p(S), _p(SQ)))
        return img, S, SQ

    def level_tilted(self, idx, lw, lh):
        """Tilted integral plane of a level (only for cascades with tilted features)."""
        T = np.zeros((lh + 1, lw + 1), np.int32)
        self._ck(lib().ugd_get_level_tilted(self._h, idx, _p(T)))
        return T


def _rects_np(buf, n):
    if n == 0:
        return np.zeros((0, 4), np.int32)
    return np.frombuffer(buf, dtype=np.int32, count=4 * n).reshape(n, 4).copy()


class FaceDetector:
    """Mirror of ugproj::FaceDetector (reference src/tracker/detector.h:10-19).

    FaceDetector(cascade, cfg) holds references to a loaded cascade and the detection config;
    detectFaces(frame, rects) appends (x, y, w, h) tuples to `rects` (detector.cc:41)."""

    def __init__(self, cascade: Cascade, cfg: Params | None = None, max_size=(1920, 1080), device: int = 0):
        self.cascade = cascade
        self.cfg_ = cfg if cfg is not None else default_params()
        self._det = Detector(cascade, self.cfg_, device=device, max_size=max_size, max_batch=1)

    def detectFaces(self, frame: np.ndarray, rects: list) -> None:
        for r in self._det.detect_bgr(frame):
            rects.append(tuple(int(v) for v in r))
