"""T0: the oracle's OpenCV / libc restatements vs the real cv2 4.13 and glibc of this image (CPU only)."""
import ctypes as C

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

LEVELS_VGA = [(640, 480), (533, 400), (444, 333), (370, 278), (309, 231), (257, 193), (214, 161), (179, 134)]
LEVELS_HD = [(1920, 1080), (1600, 900), (1333, 750), (1111, 625)]


def _img(rng, w, h, smooth):
    img = rng.integers(0, 256, (h, w), dtype=np.uint8)
    return cv2.GaussianBlur(img, (5, 5), 1.2) if smooth else img


@pytest.mark.parametrize("smooth", [False, True])
def test_resize_chain_levels(oracle, smooth):
    rng = np.random.default_rng(3)
    for chain in (LEVELS_VGA, LEVELS_HD):
        for (sw, sh), (dw, dh) in zip(chain[:-1], chain[1:]):
            img = _img(rng, sw, sh, smooth)
            assert (cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR) == oracle.resize_linear(img, dw, dh)).all()


def test_border_reflect101(oracle):
    img = np.random.default_rng(4).integers(0, 256, (57, 83), dtype=np.uint8)
    assert (cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101) == oracle.border101(img, 19)).all()


def test_gauss7(oracle):
    rng = np.random.default_rng(5)
    for w, h in [(640, 480), (179, 134), (37, 38), (8, 9)]:
        img = rng.integers(0, 256, (h, w), dtype=np.uint8)
        assert (cv2.GaussianBlur(img, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101) == oracle.gauss7(img)).all()


def test_fast9(oracle):
    rng = np.random.default_rng(6)
    for trial in range(40):
        w, h = int(rng.integers(7, 70)), int(rng.integers(7, 70))
        img = _img(rng, w, h, trial % 2)
        for th in (7, 20):
            for nm in (True, False):
                k = cv2.FastFeatureDetector_create(th, nm).detect(img, None)
                a = np.array([(p.pt[0], p.pt[1], p.response) for p in k], np.float32).reshape(-1, 3)
                b = oracle.fast9(img, th, nm).astype(np.float32)
                assert a.shape == b.shape and (a == b).all()


def test_fast_atan2(oracle):
    rng = np.random.default_rng(7)
    y = rng.integers(-200000, 200000, 100000).astype(np.float32)
    x = rng.integers(-200000, 200000, 100000).astype(np.float32)
    y[:10] = 0
    x[5:15] = 0
    ref = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
    assert (ref == oracle.fast_atan2_v(y, x)).all()


def test_sincosf_vs_glibc(oracle):
    tot = C.c_long()
    bad = oracle.lib().orc_sincos_mismatches(C.c_float(0.0), C.c_float(6.2832), 389, C.byref(tot))
    assert tot.value > 2_000_000 and bad == 0


def test_rand_vs_glibc(oracle):
    libc = C.CDLL("libc.so.6")
    for seed in (42, 1, 0, 123456789):
        libc.srand(seed)
        r = oracle.Rng(seed)
        assert all(libc.rand() == r.next() for _ in range(20000))


def test_knn2_tie_rule(oracle):
    rng = np.random.default_rng(8)
    q = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (400, 32), dtype=np.uint8)
    t[10] = t[20] = t[5]
    q[5] = t[10]
    t[100:110] = t[100]  # many exact ties
    q[7] = t[100]
    kn = cv2.BFMatcher_create(cv2.NORM_HAMMING).knnMatch(q, t, 2)
    ia = np.array([[m[0].trainIdx, m[1].trainIdx] for m in kn])
    da = np.array([[m[0].distance, m[1].distance] for m in kn])
    ib, db = oracle.knn2(q, t)
    assert (ia == ib).all() and (da == db).all()


def test_orb_tables(oracle):
    t = oracle.Orb(1000).tables()
    assert list(t["quota"]) == [217, 181, 151, 126, 105, 87, 73, 60]
    assert list(t["umax"]) == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert list(oracle.Orb(2000).tables()["quota"]) == [434, 362, 302, 251, 209, 175, 145, 122]
    assert list(oracle.Orb(4000).tables()["quota"]) == [869, 724, 603, 503, 419, 349, 291, 242]


def test_orb_pyramid_vs_cv2(oracle, seq_vga):
    """The oracle's pyramid (chain resize + 19 px REFLECT_101 ring) and blurred levels equal cv2's."""
    cam, frames = seq_vga
    orb = oracle.Orb(1000)
    orb.extract(frames[0][0])
    prev = frames[0][0]
    for l in range(8):
        w, h = orb.level_size(l)
        cur = prev if l == 0 else cv2.resize(prev, (w, h), interpolation=cv2.INTER_LINEAR)
        assert (cv2.copyMakeBorder(cur, 19, 19, 19, 19, cv2.BORDER_REFLECT_101) == orb.level_padded(l)).all()
        assert (cv2.GaussianBlur(cur, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101) == orb.level_blurred(l)).all()
        prev = cur


def test_label_quadtree_formulation_matches_list_restatement(oracle, seq_vga):
    """The pass-synchronous label formulation (what k_quadtree runs) == the std::list restatement."""
    cam, frames = seq_vga
    L = oracle.lib()
    for nf in (1000, 300):
        orb = oracle.Orb(nf)
        q = orb.tables()["quota"]
        orb.extract(frames[0][0])
        for l in range(8):
            c = orb.level_candidates(l)
            w, h = orb.level_size(l)
            packed = (c["x"].astype(np.uint32) | (c["y"].astype(np.uint32) << 12) | (c["response"].astype(np.uint32) << 24)).astype(np.uint32)
            for N in (int(q[l]), 5, 37, len(c) + 10):
                ref = oracle.quadtree(c, 16, w - 16, 16, h - 16, N)
                out = np.zeros(len(c) + 8, np.uint32)
                n = L.orc_quadtree_labels(packed.ctypes.data_as(C.c_void_p), len(c), 16, w - 16, 16, h - 16, N, out.ctypes.data_as(C.c_void_p))
                rp = ref["x"].astype(np.uint32) | (ref["y"].astype(np.uint32) << 12) | (ref["response"].astype(np.uint32) << 24)
                assert n == len(ref) and (out[:n] == rp).all()
