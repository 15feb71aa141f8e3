"""ctypes binding of librgbdfe.so (include/rgbdfe.h) for the tests and bench.py.

Plumbing only: every call goes straight through the C ABI; there is no Python or
CPU implementation behind it.  Importing works without a GPU (the library loads
and its symbols can be inspected); creating a Context without a CUDA device
raises RgbdfeError.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "librgbdfe.so")

KP_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")]
)
DM_DTYPE = np.dtype([("queryIdx", "<i4"), ("trainIdx", "<i4"), ("imgIdx", "<i4"), ("distance", "<f4")])
SE3_DTYPE = np.dtype(
    [("ok", "<i4"), ("rmse", "<f4"), ("T21", "<f4", (16,)), ("n_inliers", "<i4"), ("iters_executed", "<i4"), ("n_matches", "<i4")]
)
assert KP_DTYPE.itemsize == 28 and DM_DTYPE.itemsize == 16 and SE3_DTYPE.itemsize == 84


class RgbdfeError(RuntimeError):
    pass


_lib = None


class UmeParams(C.Structure):  # rgbdfe_umeyama_params
    _fields_ = [("errorVersion", C.c_int), ("inlierThresholdEuclidean", C.c_double), ("minimalInlierRatioThreshold", C.c_double),
                ("minimalNumberOfMatches", C.c_int), ("usedPairs", C.c_int)]


class VerifyParams(C.Structure):  # rgbdfe_verify_params
    _fields_ = [("ratio", C.c_float), ("min_matches", C.c_int), ("iterations", C.c_int), ("min_inlier_th", C.c_uint),
                ("max_mahal", C.c_float), ("sample_size", C.c_uint), ("seed", C.c_uint)]


class UmeResult(C.Structure):  # rgbdfe_umeyama_result
    _fields_ = [("ok", C.c_int), ("T", C.c_float * 16), ("n_inliers", C.c_int), ("iters_executed", C.c_int)]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RgbdfeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)"
            )
        _lib = C.CDLL(LIB_PATH)
        _lib.rgbdfe_last_error.restype = C.c_char_p
        _lib.rgbdfe_stream.restype = C.c_void_p
        _lib.rgbdfe_launch_count.restype = C.c_longlong
        for name in ("rgbdfe_last_error", "rgbdfe_stream", "rgbdfe_launch_count", "rgbdfe_destroy", "rgbdfe_sync"):
            getattr(_lib, name).argtypes = [C.c_void_p]
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Frame:
    def __init__(self, ctx):
        self.ctx = ctx
        h = C.c_void_p()
        ctx._chk(lib().rgbdfe_frame_create(ctx.h, C.byref(h)))
        self.h = h

    def download(self):
        cap = self.ctx.kp_cap
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        xyz = np.zeros((cap, 3), np.float32)
        outl = np.zeros(cap, np.uint8)
        n = C.c_int()
        self.ctx._chk(lib().rgbdfe_frame_download(self.ctx.h, self.h, _p(kps), _p(desc), _p(xyz), _p(outl), cap, C.byref(n)))
        n = n.value
        return kps[:n].copy(), desc[:n].copy(), xyz[:n].copy(), outl[:n].copy()

    def download_undistorted(self):
        pts = np.zeros((self.ctx.kp_cap, 2), np.float32)
        cell = np.zeros(self.ctx.kp_cap, np.int32)
        n = C.c_int()
        self.ctx._chk(lib().rgbdfe_frame_download_undistorted(self.ctx.h, self.h, _p(pts), _p(cell), self.ctx.kp_cap, C.byref(n)))
        return pts[: n.value].copy(), cell[: n.value].copy()

    def close(self):
        if self.h:
            lib().rgbdfe_frame_destroy(self.ctx.h, self.h)
            self.h = None


PART_NONE, PART_THROUGHPUT, PART_LATENCY = 0, 1, 2


def device_partition(device, latency_sms):
    """Split the device's SMs (CUDA green contexts) into a latency and a throughput partition.
    Returns (latency_sms, throughput_sms) as provisioned, or None when the driver has no green contexts."""
    a, b = C.c_int(0), C.c_int(0)
    rc = lib().rgbdfe_device_partition(int(device), int(latency_sms), C.byref(a), C.byref(b))
    if rc != 0:
        return None
    return a.value, b.value


class Camera(C.Structure):
    _fields_ = [("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("dist", C.c_float * 5),
                ("depth_map_factor", C.c_float)]


class Projection(C.Structure):
    """rgbdfe_projection: what Matcher::projectionMatch reads besides the keypoint arrays."""
    _fields_ = [("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("Rwc1", C.c_float * 9), ("Ow1", C.c_float * 3),
                ("Rcw2", C.c_float * 9), ("tcw2", C.c_float * 3), ("min_x", C.c_float), ("max_x", C.c_float), ("min_y", C.c_float),
                ("max_y", C.c_float), ("grid_w_inv", C.c_float), ("grid_h_inv", C.c_float), ("th", C.c_float)]

    @classmethod
    def make(cls, proj, th=5.0):
        """proj: dict(fx fy cx cy, Rwc1, Ow1, Rcw2, tcw2, bounds=(minX, maxX, minY, maxY), inv=(wInv, hInv))."""
        f = lambda v, k: (C.c_float * k)(*[float(x) for x in np.asarray(v, np.float32).reshape(-1)])
        b, i = proj["bounds"], proj["inv"]
        return cls(proj["fx"], proj["fy"], proj["cx"], proj["cy"], f(proj["Rwc1"], 9), f(proj["Ow1"], 3), f(proj["Rcw2"], 9), f(proj["tcw2"], 3),
                   float(b[0]), float(b[1]), float(b[2]), float(b[3]), float(i[0]), float(i[1]), float(th))


def camera_bounds(cam, w, h):
    """Frame::computeImageBounds + grid cell sizes -> ((minX, maxX, minY, maxY), (wInv, hInv)) as float32."""
    c = Camera(cam["fx"], cam["fy"], cam["cx"], cam["cy"], (C.c_float * 5)(*[float(v) for v in cam.get("dist", [0.0] * 5)]), 1.0)
    b, g = (C.c_float * 4)(), (C.c_float * 2)()
    rc = lib().rgbdfe_camera_bounds(C.byref(c), w, h, b, g)
    if rc != 0:
        raise RgbdfeError(f"rgbdfe_camera_bounds: {rc}")
    return tuple(np.float32(v) for v in b), tuple(np.float32(v) for v in g)


class Context:
    def __init__(self, device=0, max_w=640, max_h=480, max_feats=1000, max_batch=1):
        self.h = C.c_void_p()
        rc = lib().rgbdfe_create(C.byref(self.h), device, max_w, max_h, max_feats, max_batch)
        if rc != 0:
            self.h = None
            raise RgbdfeError(f"rgbdfe_create failed ({rc}): no usable CUDA device or bad arguments; there is no CPU fallback")
        self.max_feats = max_feats
        self.kp_cap = (max_feats + 64 + 31) // 32 * 32
        self.max_batch = max_batch
        self.nlevels = 8

    def _chk(self, rc):
        if rc != 0:
            raise RgbdfeError(f"rgbdfe error {rc}: {lib().rgbdfe_last_error(self.h).decode()}")

    def close(self):
        if self.h:
            lib().rgbdfe_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def stream(self):
        return lib().rgbdfe_stream(self.h)

    @property
    def launches(self):
        return lib().rgbdfe_launch_count(self.h)

    def stream_wait(self, signaler):
        self._chk(lib().rgbdfe_stream_wait(self.h, signaler.h))

    def set_stream_priority(self, high=True):
        self._chk(lib().rgbdfe_set_stream_priority(self.h, int(high)))

    def bind_partition(self, role):
        """role: PART_NONE / PART_THROUGHPUT / PART_LATENCY (after device_partition() split the device)."""
        self._chk(lib().rgbdfe_bind_partition(self.h, int(role)))

    def share_stream(self, owner):
        self._chk(lib().rgbdfe_share_stream(self.h, owner.h))

    def set_option(self, name, value):
        self._chk(lib().rgbdfe_set_option(self.h, name.encode(), int(value)))

    def sync(self):
        self._chk(lib().rgbdfe_sync(self.h))

    # ---- Extractor / ORBextractor ----
    def orb_configure(self, nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7):
        self._chk(lib().rgbdfe_orb_configure(self.h, nfeatures, C.c_float(scale), nlevels, ini_th, min_th))
        self.nlevels = nlevels

    def orb_tables(self):
        n = self.nlevels
        f = [np.zeros(n, np.float32) for _ in range(4)]
        q = np.zeros(n, np.int32)
        self._chk(lib().rgbdfe_orb_tables(self.h, _p(f[0]), _p(f[1]), _p(f[2]), _p(f[3]), _p(q)))
        return dict(scale=f[0], inv_scale=f[1], sigma2=f[2], inv_sigma2=f[3], quota=q)

    def orb_extract(self, gray):
        gray = np.ascontiguousarray(gray, np.uint8)
        cap = self.kp_cap
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int()
        self._chk(lib().rgbdfe_orb_extract(self.h, _p(gray), gray.shape[1], gray.shape[0], gray.strides[0], _p(kps), _p(desc), cap, C.byref(n)))
        return kps[: n.value].copy(), desc[: n.value].copy()

    def pyramid_level(self, level):
        w, h = C.c_int(), C.c_int()
        self._chk(lib().rgbdfe_orb_pyramid_level(self.h, level, None, 0, C.byref(w), C.byref(h)))
        out = np.zeros((h.value + 38, w.value + 38), np.uint8)
        self._chk(lib().rgbdfe_orb_pyramid_level(self.h, level, _p(out), out.strides[0], C.byref(w), C.byref(h)))
        return out

    def stage_blurred(self, level):
        w, h = C.c_int(), C.c_int()
        self._chk(lib().rgbdfe_orb_pyramid_level(self.h, level, None, 0, C.byref(w), C.byref(h)))
        out = np.zeros((h.value, w.value), np.uint8)
        self._chk(lib().rgbdfe_orb_stage_blurred(self.h, level, _p(out), out.strides[0]))
        return out

    def stage_candidates(self, level):
        cap = 1 << 17
        out = np.zeros(cap, KP_DTYPE)
        n = C.c_int()
        self._chk(lib().rgbdfe_orb_stage_candidates(self.h, level, _p(out), cap, C.byref(n)))
        return out[: n.value].copy()

    # ---- Extractor(ORB, ORB): the cv::ORB branch (row N3) ----
    def cvorb_configure(self, nfeatures=1000, scale=1.2, nlevels=8):
        """Extractor(ORB, ORB).setParameters: the extraction entry points now run cv::ORB (Extractor.cpp:55-59,163-166)."""
        self._chk(lib().rgbdfe_cvorb_configure(self.h, nfeatures, C.c_float(scale), nlevels))
        self.nlevels = nlevels

    def cvorb_stage_level(self, level, which=0):
        w, h = C.c_int(), C.c_int()
        self._chk(lib().rgbdfe_cvorb_stage_level(self.h, level, which, None, 0, C.byref(w), C.byref(h)))
        out = np.zeros((h.value, w.value), np.uint8)
        self._chk(lib().rgbdfe_cvorb_stage_level(self.h, level, which, _p(out), out.strides[0], C.byref(w), C.byref(h)))
        return out

    def cvorb_stage_harris(self, level):
        cap = 1 << 16
        out = np.zeros(cap, KP_DTYPE)
        n = C.c_int()
        self._chk(lib().rgbdfe_cvorb_stage_harris(self.h, level, _p(out), cap, C.byref(n)))
        return out[: n.value].copy()

    # ---- Matcher ----
    def match(self, qdesc, tdesc, q_outlier, q_xyz, t_xyz, ratio=0.9, discard_outliers=True):
        qdesc = np.ascontiguousarray(qdesc, np.uint8)
        tdesc = np.ascontiguousarray(tdesc, np.uint8)
        q_xyz = np.ascontiguousarray(q_xyz, np.float32)
        t_xyz = np.ascontiguousarray(t_xyz, np.float32)
        if q_outlier is not None:
            q_outlier = np.ascontiguousarray(q_outlier, np.uint8)
        out = np.zeros(max(len(qdesc), 1), DM_DTYPE)
        n = C.c_int()
        self._chk(
            lib().rgbdfe_match_knn2(
                self.h, _p(qdesc), len(qdesc), _p(tdesc), len(tdesc), _p(q_outlier), _p(q_xyz), _p(t_xyz), C.c_float(ratio),
                int(discard_outliers), _p(out), len(out), C.byref(n),
            )
        )
        return out[: n.value].copy()

    def projection_match(self, q_desc, q_xyz, q_outlier, t_desc, t_xyz, t_pts_un, proj, th=5.0):
        """Matcher::projectionMatch on host arrays (Features/Matcher.cpp:35-104)."""
        q_desc = np.ascontiguousarray(q_desc, np.uint8)
        t_desc = np.ascontiguousarray(t_desc, np.uint8)
        q_xyz = np.ascontiguousarray(q_xyz, np.float32)
        t_xyz = np.ascontiguousarray(t_xyz, np.float32)
        t_pts_un = np.ascontiguousarray(t_pts_un, np.float32)
        if q_outlier is not None:
            q_outlier = np.ascontiguousarray(q_outlier, np.uint8)
        P = proj if isinstance(proj, Projection) else Projection.make(proj, th)
        out = np.zeros(max(len(q_desc), 1), DM_DTYPE)
        n = C.c_int()
        self._chk(lib().rgbdfe_projection_match(self.h, _p(q_desc), len(q_desc), _p(q_xyz), _p(q_outlier), _p(t_desc), len(t_desc), _p(t_xyz),
                                                _p(t_pts_un), C.byref(P), _p(out), len(out), C.byref(n)))
        return out[: n.value].copy()

    def frames_projection_match(self, f1, f2, projs):
        """projectionMatch for device-resident frame pairs; projs: one Projection (or dict) per pair.  Returns a list of match arrays."""
        npairs = len(f1)
        P = (Projection * npairs)(*[p if isinstance(p, Projection) else Projection.make(p, p.get("th", 5.0)) for p in projs])
        out = np.empty((npairs, self.kp_cap), DM_DTYPE)
        n = np.zeros(npairs, np.int32)
        self._chk(lib().rgbdfe_frames_projection_match(self.h, (C.c_void_p * npairs)(*[f.h for f in f1]), (C.c_void_p * npairs)(*[f.h for f in f2]),
                                                       P, npairs, _p(out), self.kp_cap, _p(n)))
        return [out[i, : n[i]].copy() for i in range(npairs)]

    def projection_rounds(self, npairs=1):
        r = np.zeros(npairs, np.int32)
        self._chk(lib().rgbdfe_projection_rounds(self.h, _p(r), npairs))
        return r

    # ---- Random ----
    def rng_seed(self, seed):
        self._chk(lib().rgbdfe_rng_seed(self.h, C.c_uint(seed)))

    def rng_next(self):
        v = C.c_int()
        self._chk(lib().rgbdfe_rng_next(self.h, C.byref(v)))
        return v.value

    def rng_get_state(self):
        st = np.zeros(35, np.int32)
        self._chk(lib().rgbdfe_rng_get_state(self.h, _p(st)))
        return st

    def rng_set_state(self, st):
        st = np.ascontiguousarray(st, np.int32)
        self._chk(lib().rgbdfe_rng_set_state(self.h, _p(st)))

    def depth_cov_get(self):
        l, cov = C.c_int(), C.c_double()
        self._chk(lib().rgbdfe_depth_cov_get(self.h, C.byref(l), C.byref(cov)))
        return l.value, cov.value

    def depth_cov_set(self, latched, cov):
        self._chk(lib().rgbdfe_depth_cov_set(self.h, int(latched), C.c_double(cov)))

    # ---- RansacSE3 ----
    def ransac_se3(self, p1, p2, matches, iters=200, min_inl=10, max_mahal=3.0, sample=4):
        p1 = np.ascontiguousarray(p1, np.float32)
        p2 = np.ascontiguousarray(p2, np.float32)
        matches = np.ascontiguousarray(matches, DM_DTYPE)
        res = np.zeros(1, SE3_DTYPE)
        inl = np.zeros(max(len(matches), 1), DM_DTYPE)
        self._chk(
            lib().rgbdfe_ransac_se3(
                self.h, _p(p1), len(p1), _p(p2), len(p2), _p(matches), len(matches), iters, C.c_uint(min_inl), C.c_float(max_mahal),
                C.c_uint(sample), _p(res), _p(inl), len(inl),
            )
        )
        r = res[0]
        return dict(
            ok=bool(r["ok"]), rmse=np.float32(r["rmse"]), T=r["T21"].reshape(4, 4).copy(), inliers=inl[: r["n_inliers"]].copy(),
            iters_executed=int(r["iters_executed"]), n_matches=int(r["n_matches"]),
        )

    # ---- Ransac (Umeyama) ----
    def ransac_umeyama(self, p1, p2, matches, error_version=0, th=0.02, min_ratio=0.2, min_matches=10, used_pairs=3):
        p1 = np.ascontiguousarray(p1, np.float32)
        p2 = np.ascontiguousarray(p2, np.float32)
        matches = np.ascontiguousarray(matches, DM_DTYPE)
        prm = UmeParams(error_version, th, min_ratio, min_matches, used_pairs)
        res = UmeResult()
        inl = np.zeros(max(len(matches), 1), DM_DTYPE)
        self._chk(
            lib().rgbdfe_ransac_umeyama(
                self.h, _p(p1), len(p1), _p(p2), len(p2), _p(matches), len(matches), C.byref(prm), C.byref(res), _p(inl), len(inl)
            )
        )
        return dict(ok=bool(res.ok), T=np.array(res.T, np.float32).reshape(4, 4), inliers=inl[: res.n_inliers].copy(),
                    iters_executed=int(res.iters_executed))

    # ---- device-resident pipeline ----
    def frames_extract(self, frames, grays, depths, cam):
        n = len(frames)
        fr = (C.c_void_p * n)(*[f.h for f in frames])
        gp = (C.c_void_p * n)(*[g.ctypes.data for g in grays])
        h, w = grays[0].shape
        if depths is not None:
            dp = (C.c_void_p * n)(*[d.ctypes.data for d in depths])
            dstride = depths[0].strides[0] // 4
        else:
            dp, dstride = None, 0
        self._chk(
            lib().rgbdfe_frames_extract(
                self.h, fr, n, gp, w, h, grays[0].strides[0], dp, dstride, C.c_float(cam["fx"]), C.c_float(cam["fy"]), C.c_float(cam["cx"]),
                C.c_float(cam["cy"]),
            )
        )

    def frames_extract_raw(self, frames, grays, depths16, cam, depth_map_factor):
        """gray u8 + raw 16-bit depth; depth_map_factor = 1.0f / DepthMapFactor (Core/RGBDcamera.cpp:21)."""
        n = len(frames)
        fr = (C.c_void_p * n)(*[f.h for f in frames])
        gp = (C.c_void_p * n)(*[g.ctypes.data for g in grays])
        dp = (C.c_void_p * n)(*[d.ctypes.data for d in depths16])
        h, w = grays[0].shape
        assert depths16[0].dtype == np.uint16
        self._chk(
            lib().rgbdfe_frames_extract_raw(
                self.h, fr, n, gp, w, h, grays[0].strides[0], dp, depths16[0].strides[0] // 2, C.c_float(depth_map_factor),
                C.c_float(cam["fx"]), C.c_float(cam["fy"]), C.c_float(cam["cx"]), C.c_float(cam["cy"]),
            )
        )

    def frames_from_images(self, frames, images, depths16, cam):
        """N1: images are gray (H, W) or BGR (H, W, 3) uint8; cam: fx fy cx cy dist[5] factor (DepthMapFactor)."""
        n = len(frames)
        fr = (C.c_void_p * n)(*[f.h for f in frames])
        ip = (C.c_void_p * n)(*[g.ctypes.data for g in images])
        dp = (C.c_void_p * n)(*[d.ctypes.data for d in depths16])
        h, w = images[0].shape[:2]
        ch = 1 if images[0].ndim == 2 else images[0].shape[2]
        cs = Camera(cam["fx"], cam["fy"], cam["cx"], cam["cy"], (C.c_float * 5)(*[float(v) for v in cam.get("dist", [0] * 5)]),
                    float(np.float32(1.0) / np.float32(cam["factor"])))
        self._chk(lib().rgbdfe_frames_from_images(self.h, fr, n, ip, ch, w, h, images[0].strides[0], dp, depths16[0].strides[0] // 2, C.byref(cs)))

    def track(self, ref, cur, slot, ratio=0.9, discard=True, iters=200, min_inl=10, max_mahal=3.0, sample=4, update_f2=True):
        self._chk(
            lib().rgbdfe_track(
                self.h, ref.h, cur.h, C.c_float(ratio), int(discard), iters, C.c_uint(min_inl), C.c_float(max_mahal), C.c_uint(sample),
                int(update_f2), slot,
            )
        )

    def track_sequence(self, frames, first_slot, ratio=0.9, discard=True, iters=200, min_inl=10, max_mahal=3.0, sample=4, update_f2=True):
        n = len(frames)
        fr = (C.c_void_p * n)(*[f.h for f in frames])
        self._chk(
            lib().rgbdfe_track_sequence(
                self.h, fr, n, C.c_float(ratio), int(discard), iters, C.c_uint(min_inl), C.c_float(max_mahal), C.c_uint(sample), int(update_f2),
                first_slot,
            )
        )

    def track_results(self, first, count):
        res = np.zeros(count, SE3_DTYPE)
        self._chk(lib().rgbdfe_track_results(self.h, first, count, _p(res)))
        return res

    def track_results_inliers(self, first, count):
        """(results, [mvInliers of every pair]) of the last track / track_sequence call."""
        res = np.zeros(count, SE3_DTYPE)
        inl = np.zeros(count * self.kp_cap, DM_DTYPE)
        off = np.zeros(count + 1, np.int32)
        self._chk(lib().rgbdfe_track_results_inliers(self.h, first, count, _p(res), _p(inl), len(inl), _p(off)))
        return res, [inl[off[i]:off[i + 1]].copy() for i in range(count)]

    def frames_download_async(self, frames, kps, desc, xyz, n, cap):
        """kps/desc/xyz/n: numpy arrays (ideally pinned) of [len(frames)][cap] rows; complete after sync()."""
        fr = (C.c_void_p * len(frames))(*[f.h for f in frames])
        self._chk(lib().rgbdfe_frames_download_async(self.h, fr, len(frames), _p(kps), _p(desc), _p(xyz), cap, _p(n)))

    # ---- LoopDetector (row N4) ----
    def bowdb_create(self, max_kf, max_words_per_kf):
        h = C.c_void_p()
        self._chk(lib().rgbdfe_bowdb_create(self.h, C.byref(h), max_kf, max_words_per_kf))
        return h

    def bowdb_destroy(self, db):
        lib().rgbdfe_bowdb_destroy(self.h, db)

    def bowdb_add(self, db, kf_id, words, weights):
        words = np.ascontiguousarray(words, np.int32)
        weights = np.ascontiguousarray(weights, np.float64)
        idx = C.c_int()
        self._chk(lib().rgbdfe_bowdb_add(self.h, db, int(kf_id), _p(words), _p(weights), len(words), C.byref(idx)))
        return idx.value

    def loop_candidates(self, db, query_index, connected, interval):
        conn = np.ascontiguousarray(connected, np.int32)
        cand, sc, n = np.zeros(8, np.int32), np.zeros(8, np.float64), C.c_int()
        self._chk(lib().rgbdfe_loop_candidates(self.h, db, int(query_index), _p(conn), len(conn), int(interval), _p(cand), _p(sc), 8, C.byref(n)))
        return [(int(cand[i]), float(sc[i])) for i in range(n.value)]

    # ---- batched verification ----
    def kfdb_create(self, desc, xyz):
        desc = np.ascontiguousarray(desc, np.uint8)
        xyz = np.ascontiguousarray(xyz, np.float32)
        h = C.c_void_p()
        self._chk(lib().rgbdfe_kfdb_create(self.h, C.byref(h), desc.shape[0], desc.shape[1], _p(desc), _p(xyz)))
        return h

    def kfdb_destroy(self, db):
        lib().rgbdfe_kfdb_destroy(self.h, db)

    def verify_batch(self, db, n_kf, q_desc, q_xyz, ratio=0.9, min_matches=30, iters=200, min_inl=30, max_mahal=3.0, sample=4, seed=42,
                     first_pair_index=0, results_dev=None, want_host=True):
        q_desc = np.ascontiguousarray(q_desc, np.uint8)
        q_xyz = np.ascontiguousarray(q_xyz, np.float32)
        res = np.zeros(n_kf, SE3_DTYPE) if want_host else None
        self._chk(
            lib().rgbdfe_verify_batch(
                self.h, db, _p(q_desc), _p(q_xyz), len(q_desc), C.c_float(ratio), min_matches, iters, C.c_uint(min_inl), C.c_float(max_mahal),
                C.c_uint(sample), C.c_uint(seed), first_pair_index, _p(res), C.c_void_p(results_dev) if results_dev else None,
            )
        )
        return res

    # ---- sharded verification (rgbdfe_verify_sharded_*; NCCL communicator created through the C ABI) ----
    def nccl_comm_create(self, n_ranks, rank, id128):
        comm = C.c_void_p()
        self._chk(lib().rgbdfe_nccl_comm_create(self.h, n_ranks, rank, id128, C.byref(comm)))
        return comm

    def verify_sharded_enqueue(self, db, comm, n_ranks, rank, n_kf_total, q_desc, q_xyz, nq, n_queries, params, slot):
        """q_desc / q_xyz: numpy arrays (host) or integer device addresses."""
        pd = _p(q_desc) if isinstance(q_desc, np.ndarray) else C.c_void_p(int(q_desc))
        px = _p(q_xyz) if isinstance(q_xyz, np.ndarray) else C.c_void_p(int(q_xyz))
        self._chk(lib().rgbdfe_verify_sharded_enqueue(self.h, db, comm, n_ranks, rank, n_kf_total, pd, px, nq, n_queries, C.byref(params), slot))

    def verify_sharded_collect(self, slot, n_queries, n_kf_total, want=True, out=None):
        """out: a preallocated (n_queries, n_kf_total) SE3_DTYPE array to fill (callers in a loop avoid a fresh allocation per call)."""
        res = out if out is not None else (np.zeros((n_queries, n_kf_total), SE3_DTYPE) if want else None)
        self._chk(lib().rgbdfe_verify_sharded_collect(self.h, slot, _p(res)))
        return res


def nccl_unique_id():
    buf = C.create_string_buffer(128)
    rc = lib().rgbdfe_nccl_unique_id(buf)
    if rc != 0:
        raise RgbdfeError(f"rgbdfe_nccl_unique_id: {rc} (NCCL library not found?)")
    return buf.raw


def shard_range(n_kf_total, rank, n_ranks):
    lo, hi = C.c_int(), C.c_int()
    lib().rgbdfe_shard_range.restype = None
    lib().rgbdfe_shard_range(n_kf_total, rank, n_ranks, C.byref(lo), C.byref(hi))
    return lo.value, hi.value


def gathered_to_keyframe_order(gathered, n_ranks, n_queries, n_kf_total):
    """gathered: SE3_DTYPE records in the all-gather layout [rank][query][ceil(n_kf_total / n_ranks)]."""
    g = np.ascontiguousarray(gathered)
    out = np.zeros((n_queries, n_kf_total), SE3_DTYPE)
    rc = lib().rgbdfe_gathered_to_keyframe_order(_p(g), n_ranks, n_queries, n_kf_total, _p(out))
    if rc != 0:
        raise RgbdfeError(f"rgbdfe_gathered_to_keyframe_order: {rc}")
    return out
