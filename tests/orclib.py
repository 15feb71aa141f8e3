"""ctypes loader for the CPU oracle (oracle/_build/liboracle.so).

Test infrastructure only: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product never loads it.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB_PATH = os.path.join(ORACLE_DIR, "_build", "liboracle.so")

KP_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")]
)
DM_DTYPE = np.dtype([("queryIdx", "<i4"), ("trainIdx", "<i4"), ("imgIdx", "<i4"), ("distance", "<f4")])
assert KP_DTYPE.itemsize == 28 and DM_DTYPE.itemsize == 16


class Se3Out(C.Structure):
    _fields_ = [
        ("ok", C.c_int),
        ("rmse", C.c_float),
        ("T", C.c_float * 16),
        ("n_inliers", C.c_int),
        ("iters_executed", C.c_int),
        ("es", C.c_long),
        ("ef", C.c_long),
        ("adds", C.c_long),
        ("svds", C.c_long),
        ("rounds", C.c_long),
    ]


class UmeOut(C.Structure):
    _fields_ = [("ok", C.c_int), ("T", C.c_float * 16), ("n_inliers", C.c_int), ("iters_executed", C.c_int)]


_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        _lib = C.CDLL(LIB_PATH)
        L = _lib
        L.orc_fast_atan2.restype = C.c_float
        L.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.orc_sinf.restype = C.c_float
        L.orc_sinf.argtypes = [C.c_float]
        L.orc_cosf.restype = C.c_float
        L.orc_cosf.argtypes = [C.c_float]
        L.orc_cv_round.argtypes = [C.c_float]
        L.orc_sincos_mismatches.restype = C.c_long
        L.orc_sincos_mismatches.argtypes = [C.c_float, C.c_float, C.c_int, C.POINTER(C.c_long)]
        L.orc_orb_create.restype = C.c_void_p
        L.orc_orb_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orc_orb_destroy.argtypes = [C.c_void_p]
        L.orc_mahal2.restype = C.c_double
        L.orc_cvorb_create.restype = C.c_void_p
        L.orc_cvorb_create.argtypes = [C.c_int, C.c_float, C.c_int]
        L.orc_cvorb_destroy.argtypes = [C.c_void_p]
        for name in (
            "orc_orb_extract",
            "orc_orb_tables",
            "orc_orb_level_size",
            "orc_orb_level_padded",
            "orc_orb_level_blurred",
            "orc_orb_level_candidates",
            "orc_orb_level_keypoints",
        ):
            getattr(L, name).argtypes = None
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def resize_linear(src, dw, dh):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty((dh, dw), np.uint8)
    lib().orc_resize_linear_u8(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dw, dh, dw)
    return dst


def border101(src, b):
    src = np.ascontiguousarray(src, np.uint8)
    h, w = src.shape
    dst = np.empty((h + 2 * b, w + 2 * b), np.uint8)
    lib().orc_border_reflect101(_p(src), w, h, src.strides[0], _p(dst), w + 2 * b, b)
    return dst


def fast9(img, th, nonmax=True):
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    out = np.empty((w * h, 3), np.int32)
    n = lib().orc_fast9(_p(img), w, h, img.strides[0], int(th), int(nonmax), _p(out), w * h)
    return out[:n].copy()


def gauss7(img):
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    dst = np.empty_like(img)
    lib().orc_gauss7_u8(_p(img), w, h, img.strides[0], _p(dst), w)
    return dst


def fast_atan2_v(y, x):
    y = np.ascontiguousarray(y, np.float32)
    x = np.ascontiguousarray(x, np.float32)
    out = np.empty_like(y)
    lib().orc_fast_atan2_v(_p(y), _p(x), _p(out), y.size)
    return out


def knn2(q, t):
    q = np.ascontiguousarray(q, np.uint8)
    t = np.ascontiguousarray(t, np.uint8)
    idx = np.empty((len(q), 2), np.int32)
    dist = np.empty((len(q), 2), np.int32)
    lib().orc_knn2_hamming(_p(q), len(q), _p(t), len(t), _p(idx), _p(dist))
    return idx, dist


def resize_linear_exact(src, dw, dh):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty((dh, dw), np.uint8)
    lib().orc_resize_linear_exact(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dw, dh, dw)
    return dst


def gauss7_float(img):
    """the blur cv::ORB applies to its pyramid levels (float sepFilter2D path, oracle/cvorb.cpp)"""
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    dst = np.empty_like(img)
    lib().orc_gauss7_float(_p(img), w, h, img.strides[0], _p(dst), w)
    return dst


class CvOrb:
    """Extractor(ORB, ORB): cv::ORB::create(nfeatures, scale, nlevels).detect + cv::ORB::create().compute
    (Features/Extractor.cpp:55-59,163-166,216-218), restated in oracle/cvorb.cpp."""

    def __init__(self, nfeatures=1000, scale=1.2, nlevels=8):
        self.h = C.c_void_p(lib().orc_cvorb_create(nfeatures, scale, nlevels))
        self.nlevels = nlevels
        self.cap = 4 * nfeatures + 64

    def extract(self, img):
        img = np.ascontiguousarray(img, np.uint8)
        kps = np.zeros(self.cap, KP_DTYPE)
        desc = np.zeros((self.cap, 32), np.uint8)
        n = lib().orc_cvorb_extract(self.h, _p(img), img.shape[1], img.shape[0], img.strides[0], _p(kps), _p(desc), self.cap)
        assert n <= self.cap
        return kps[:n].copy(), desc[:n].copy()

    def level_info(self, level):
        v = [C.c_int() for _ in range(5)]
        lib().orc_cvorb_level_size(self.h, level, *[C.byref(x) for x in v])
        return dict(zip(("w", "h", "quota", "n_fast", "n_harris"), (x.value for x in v)))

    def level_image(self, level, blurred=False):
        i = self.level_info(level)
        out = np.empty((i["h"], i["w"]), np.uint8)
        lib().orc_cvorb_level_image(self.h, level, _p(out), int(blurred))
        return out

    def level_harris(self, level):
        i = self.level_info(level)
        out = np.zeros(max(i["n_harris"], 1), KP_DTYPE)
        n = lib().orc_cvorb_level_harris(self.h, level, _p(out), len(out))
        return out[:n].copy()

    def __del__(self):
        try:
            lib().orc_cvorb_destroy(self.h)
        except Exception:
            pass


class Rng:
    def __init__(self, seed=42):
        self.buf = np.zeros(lib().orc_rng_sizeof(), np.uint8)
        self.seed(seed)

    def seed(self, s):
        lib().orc_rng_seed(_p(self.buf), C.c_uint(s))

    def next(self):
        return lib().orc_rng_next(_p(self.buf))

    def random_int(self, a, b):
        return lib().orc_rng_random_int(_p(self.buf), a, b)

    @property
    def ptr(self):
        return _p(self.buf)


class Orb:
    def __init__(self, nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7):
        self.h = lib().orc_orb_create(nfeatures, scale, nlevels, ini_th, min_th)
        self.nlevels = nlevels
        self.nfeatures = nfeatures

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_orb_destroy(C.c_void_p(self.h))
            self.h = None

    def extract(self, img):
        img = np.ascontiguousarray(img, np.uint8)
        cap = self.nfeatures * 2 + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = lib().orc_orb_extract(C.c_void_p(self.h), _p(img), img.shape[1], img.shape[0], img.strides[0], _p(kps), _p(desc), cap)
        assert n <= cap
        return kps[:n].copy(), desc[:n].copy()

    def tables(self):
        n = self.nlevels
        f = [np.zeros(n, np.float32) for _ in range(4)]
        quota = np.zeros(n, np.int32)
        umax = np.zeros(16, np.int32)
        lib().orc_orb_tables(C.c_void_p(self.h), _p(f[0]), _p(f[1]), _p(f[2]), _p(f[3]), _p(quota), _p(umax))
        return dict(scale=f[0], inv_scale=f[1], sigma2=f[2], inv_sigma2=f[3], quota=quota, umax=umax)

    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        lib().orc_orb_level_size(C.c_void_p(self.h), level, C.byref(w), C.byref(h))
        return w.value, h.value

    def level_padded(self, level):
        w, h = self.level_size(level)
        out = np.empty((h + 38, w + 38), np.uint8)
        lib().orc_orb_level_padded(C.c_void_p(self.h), level, _p(out))
        return out

    def level_blurred(self, level):
        w, h = self.level_size(level)
        out = np.empty((h, w), np.uint8)
        ok = lib().orc_orb_level_blurred(C.c_void_p(self.h), level, _p(out))
        return out if ok else None

    def level_candidates(self, level):
        cap = 1 << 17
        out = np.zeros(cap, KP_DTYPE)
        n = lib().orc_orb_level_candidates(C.c_void_p(self.h), level, _p(out), cap)
        return out[:n].copy()

    def level_keypoints(self, level):
        cap = 1 << 15
        out = np.zeros(cap, KP_DTYPE)
        n = lib().orc_orb_level_keypoints(C.c_void_p(self.h), level, _p(out), cap)
        return out[:n].copy()


def quadtree(cands, minX, maxX, minY, maxY, N):
    cands = np.ascontiguousarray(cands, KP_DTYPE)
    out = np.zeros(max(len(cands), 1), KP_DTYPE)
    n = lib().orc_quadtree(_p(cands), len(cands), minX, maxX, minY, maxY, N, _p(out), len(out))
    return out[:n].copy()


def match(qdesc, tdesc, q_outlier, q3d, t3d, ratio=0.9, discard_outliers=True):
    qdesc = np.ascontiguousarray(qdesc, np.uint8)
    tdesc = np.ascontiguousarray(tdesc, np.uint8)
    q_outlier = np.ascontiguousarray(q_outlier, np.uint8)
    q3d = np.ascontiguousarray(q3d, np.float32)
    t3d = np.ascontiguousarray(t3d, np.float32)
    out = np.zeros(max(len(qdesc), 1), DM_DTYPE)
    n = lib().orc_match(
        _p(qdesc), len(qdesc), _p(tdesc), len(tdesc), _p(q_outlier), _p(q3d), _p(t3d), C.c_float(ratio), int(discard_outliers), _p(out), len(out)
    )
    return out[:n].copy()


def sort_matches(m):
    m = np.ascontiguousarray(m, DM_DTYPE).copy()
    lib().orc_sort_matches(_p(m), len(m))
    return m


class Latch:
    def __init__(self):
        self.latched = C.c_int(0)
        self.cov = C.c_double(0.0)


def ransac_se3(p1, p2, matches, rng, latch, iters=200, min_inl=10, max_mahal=3.0, sample=4):
    p1 = np.ascontiguousarray(p1, np.float32)
    p2 = np.ascontiguousarray(p2, np.float32)
    matches = np.ascontiguousarray(matches, DM_DTYPE)
    out = Se3Out()
    inl = np.zeros(max(len(matches), 1), DM_DTYPE)
    lib().orc_ransac_se3(
        _p(p1), _p(p2), _p(matches), len(matches), iters, C.c_uint(min_inl), C.c_float(max_mahal), C.c_uint(sample),
        rng.ptr, C.byref(latch.latched), C.byref(latch.cov), C.byref(out), _p(inl), len(inl),
    )
    return dict(
        ok=bool(out.ok), rmse=np.float32(out.rmse), T=np.array(out.T, np.float32).reshape(4, 4), inliers=inl[: out.n_inliers].copy(),
        iters_executed=out.iters_executed, es=out.es, ef=out.ef, adds=out.adds, svds=out.svds, rounds=out.rounds,
    )


def ransac_umeyama(p1, p2, matches, rng, error_version=0, th=0.02, min_ratio=0.2, min_matches=10, used_pairs=3):
    p1 = np.ascontiguousarray(p1, np.float32)
    p2 = np.ascontiguousarray(p2, np.float32)
    matches = np.ascontiguousarray(matches, DM_DTYPE)
    out = UmeOut()
    inl = np.zeros(max(len(matches), 1), DM_DTYPE)
    lib().orc_ransac_umeyama(
        _p(p1), _p(p2), _p(matches), len(matches), error_version, C.c_double(th), C.c_double(min_ratio), min_matches, used_pairs,
        rng.ptr, C.byref(out), _p(inl), len(inl),
    )
    return dict(ok=bool(out.ok), T=np.array(out.T, np.float32).reshape(4, 4), inliers=inl[: out.n_inliers].copy(), iters_executed=out.iters_executed)
