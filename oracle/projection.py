"""CPU restatement (numpy) of Matcher::projectionMatch, row N2 of the scope table -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module; the product never does.
Each function cites the reference lines it follows (paths relative to the reference tree).  The reference never
calls projectionMatch and holds no test or golden vector for it, so the pin is on the primitives: the 3x3 * 3x1
cv::Mat expression against cv2.gemm (tests/test_projection.py live, tests/golden/projection_cv2.npz), the Hamming
norm against cv2.norm, posInGrid / the image bounds through oracle/frame_prep.py (pinned there).
"""
import numpy as np

FRAME_GRID_COLS, FRAME_GRID_ROWS = 64, 48  # Core/Frame.h:15-16
TH_HIGH = 100  # Features/Matcher.cpp:9
INT_MIN = -(1 << 31)
_f = np.float32


def mat_vec_add(A, b, c):
    """cv::Mat `A * b + c` with A 3x3, b and c 3x1, all CV_32F (Core/Frame.cpp:324, Features/Matcher.cpp:52): the
    expression folds into cv::gemm(A, b, 1, c, 1), whose small-matrix path forms each row as float products summed
    left to right in float and then adds c (the double add of two floats, narrowed, equals the float add)."""
    A = np.asarray(A, _f).reshape(3, 3)
    b = np.asarray(b, _f).reshape(3)
    c = np.asarray(c, _f).reshape(3)
    out = np.empty(3, _f)
    with np.errstate(all="ignore"):
        for r in range(3):
            t = _f(_f(_f(A[r, 0] * b[0]) + _f(A[r, 1] * b[1])) + _f(A[r, 2] * b[2]))
            out[r] = _f(t + c[r])
    return out


def _cvtt(v):
    """(int) of a float on x86-64 (cvttss2si): truncation, INT_MIN for NaN and for values outside int's range."""
    v = float(v)
    if v != v or v >= 2147483648.0 or v < -2147483648.0:
        return INT_MIN
    return int(v)


def pos_in_grid(pts_un, bounds, inv):
    """Frame::posInGrid, Core/Frame.cpp:231-241, for every undistorted keypoint: (cell x, cell y) or (-1, -1)."""
    import frame_prep

    pts_un = np.asarray(pts_un, _f).reshape(-1, 2)
    gx = frame_prep._roundf((pts_un[:, 0] - _f(bounds[0])) * _f(inv[0])).astype(np.int64)
    gy = frame_prep._roundf((pts_un[:, 1] - _f(bounds[2])) * _f(inv[1])).astype(np.int64)
    ok = (gx >= 0) & (gx < FRAME_GRID_COLS) & (gy >= 0) & (gy < FRAME_GRID_ROWS)
    return np.where(ok, gx, -1), np.where(ok, gy, -1)


def build_grid(pts_un, bounds, inv):
    """Frame::assignFeaturesToGrid, Core/Frame.cpp:75-89: mGrid[x][y] lists in keypoint order."""
    gx, gy = pos_in_grid(pts_un, bounds, inv)
    grid = {}
    for i, (x, y) in enumerate(zip(gx, gy)):
        if x >= 0:
            grid.setdefault((int(x), int(y)), []).append(i)
    return grid


def features_in_area(grid, pts_un, bounds, inv, x, y, r):
    """Frame::getFeaturesInArea(x, y, r) with the default level range, Core/Frame.cpp:179-229."""
    out = []
    x, y, r = _f(x), _f(y), _f(r)
    min_x, min_y = _f(bounds[0]), _f(bounds[2])
    w_inv, h_inv = _f(inv[0]), _f(inv[1])
    with np.errstate(all="ignore"):
        c0 = max(0, _cvtt(np.floor(_f(_f(_f(x - min_x) - r) * w_inv))))
        if c0 >= FRAME_GRID_COLS:
            return out
        c1 = min(FRAME_GRID_COLS - 1, _cvtt(np.ceil(_f(_f(_f(x - min_x) + r) * w_inv))))
        if c1 < 0:
            return out
        r0 = max(0, _cvtt(np.floor(_f(_f(_f(y - min_y) - r) * h_inv))))
        if r0 >= FRAME_GRID_ROWS:
            return out
        r1 = min(FRAME_GRID_ROWS - 1, _cvtt(np.ceil(_f(_f(_f(y - min_y) + r) * h_inv))))
        if r1 < 0:
            return out
        for ix in range(c0, c1 + 1):
            for iy in range(r0, r1 + 1):
                for j in grid.get((ix, iy), ()):
                    if abs(_f(pts_un[j, 0] - x)) < r and abs(_f(pts_un[j, 1] - y)) < r:
                        out.append(j)
    return out


_POP = np.array([bin(i).count("1") for i in range(256)], np.int32)


def hamming(a, b):
    """Matcher::descriptorDistance = cv::norm(a, b, NORM_HAMMING), Features/Matcher.cpp:141-144."""
    return int(_POP[np.bitwise_xor(a, b)].sum())


def projection_match(f1, f2, proj, th=5.0):
    """Matcher::projectionMatch(F1, F2, vMatches12, th), Features/Matcher.cpp:35-104.

    f1: dict(desc [N1,32] u8, xyz [N1,3] f32, outlier [N1] bool)          -- F1->mDescriptors, mvKeys3Dc, mvbOutlier
    f2: dict(desc [N2,32] u8, xyz [N2,3] f32, pts_un [N2,2] f32)           -- F2->mDescriptors, mvKeys3Dc, mvKeysUn[].pt
    proj: dict(fx fy cx cy; Rwc1 [3,3], Ow1 [3]; Rcw2 [3,3], tcw2 [3]; bounds (minX maxX minY maxY); inv (wInv hInv))
    Returns a list of (queryIdx, trainIdx, distance)."""
    fx, fy, cx, cy = (_f(proj[k]) for k in ("fx", "fy", "cx", "cy"))
    bounds = [_f(v) for v in proj["bounds"]]
    inv = [_f(v) for v in proj["inv"]]
    th = _f(th)
    xyz1 = np.asarray(f1["xyz"], _f)
    xyz2 = np.asarray(f2["xyz"], _f)
    pts2 = np.asarray(f2["pts_un"], _f).reshape(-1, 2)
    grid = build_grid(pts2, bounds, inv)
    claimed, out = set(), []
    with np.errstate(all="ignore"):
        for i1 in range(len(xyz1)):
            if f1["outlier"] is not None and f1["outlier"][i1]:
                continue  # :46
            if not xyz1[i1, 2] > 0:
                continue  # :48 isValidObs
            xw = mat_vec_add(proj["Rwc1"], xyz1[i1], proj["Ow1"])  # :51 Frame::unprojectWorld, Core/Frame.cpp:317-327
            xc, yc, zc = mat_vec_add(proj["Rcw2"], xw, proj["tcw2"])  # :52
            invzc = _f(1.0) / zc
            if invzc < 0:
                continue  # :58
            u = _f(_f(_f(fx * xc) * invzc) + cx)
            v = _f(_f(_f(fy * yc) * invzc) + cy)
            if u < bounds[0] or u > bounds[1]:
                continue  # :64
            if v < bounds[2] or v > bounds[3]:
                continue
            best, best_idx = None, -1
            for i2 in features_in_area(grid, pts2, bounds, inv, u, v, th):
                if not xyz2[i2, 2] > 0 or i2 in claimed:
                    continue
                d = hamming(f1["desc"][i1], f2["desc"][i2])
                if best is None or d < best:
                    best, best_idx = d, i2
            if best is not None and best <= TH_HIGH:
                out.append((i1, best_idx, float(best)))
                claimed.add(best_idx)
    return out
