"""CPU restatement (numpy) of Frame's image preprocessing, row N1 of the scope table -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module; the product never does.
Each function cites the reference lines it follows (paths relative to the reference tree) and the OpenCV routine
those lines call.  Pinned bit-for-bit against cv2 4.13 (tests/test_oracle_cv2.py live, tests/golden/frame_prep_cv2.npz).
"""
import numpy as np

FRAME_GRID_COLS, FRAME_GRID_ROWS = 64, 48  # Core/Frame.h:15-16


def bgr2gray(bgr):
    """cvtColor(imRGB, mImGray, CV_BGR2GRAY), Core/Frame.cpp:47.  OpenCV 4.x 8-bit path: 15-bit fixed point
    (B 3735, G 19235, R 9798), rounded.  (OpenCV 3.x used a 14-bit table; the author's version is not pinned.)"""
    b, g, r = (bgr[..., i].astype(np.int64) for i in range(3))
    return ((b * 3735 + g * 19235 + r * 9798 + 16384) >> 15).astype(np.uint8)


def depth_to_float(d16, depth_map_factor):
    """imDepth.convertTo(mImDepth, CV_32F, mDepthMapFactor), Core/Frame.cpp:48: float(d) * float(factor), one rounding."""
    return d16.astype(np.float32) * np.float32(depth_map_factor)


def undistort_points(pts, fx, fy, cx, cy, dist):
    """cv::undistortPoints(mat, mat, mK, distCoef, cv::Mat(), mK), Core/Frame.cpp:270 (and :300 for the corners).
    OpenCV undistort.dispatch.cpp: K and distCoef (CV_32F) widened to double, five iterations, result narrowed to float."""
    fx, fy, cx, cy = (float(np.float32(v)) for v in (fx, fy, cx, cy))
    k = [float(np.float32(v)) for v in dist] + [0.0] * (12 - len(dist))
    ifx, ify = 1.0 / fx, 1.0 / fy
    out = np.empty((len(pts), 2), np.float32)
    for i, (u, v) in enumerate(np.asarray(pts, np.float32).astype(np.float64)):
        x = (u - cx) * ifx
        y = (v - cy) * ify
        x0, y0 = x, y
        for _ in range(5):
            r2 = x * x + y * y
            icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2)
            if icdist < 0:
                x, y = x0, y0
                break
            dx = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2
            dy = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2
            x = (x0 - dx) * icdist
            y = (y0 - dy) * icdist
        out[i] = (np.float32(fx * x + cx), np.float32(fy * y + cy))
    return out


def undistort_keypoints(pts, cam):
    """Frame::undistortKeyPoints, Core/Frame.cpp:251-281: identity when distCoef[0] == 0."""
    pts = np.asarray(pts, np.float32).reshape(-1, 2)
    if np.float32(cam["dist"][0]) == 0:
        return pts.copy()
    return undistort_points(pts, cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["dist"])


def image_bounds(w, h, cam):
    """Frame::computeImageBounds, Core/Frame.cpp:283-316 -> (mnMinX, mnMaxX, mnMinY, mnMaxY) as float32."""
    if np.float32(cam["dist"][0]) == 0:
        return np.float32(0), np.float32(w), np.float32(0), np.float32(h)
    c = undistort_points([[0, 0], [w, 0], [0, h], [w, h]], cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["dist"])
    return min(c[0, 0], c[2, 0]), max(c[1, 0], c[3, 0]), min(c[0, 1], c[1, 1]), max(c[2, 1], c[3, 1])


def _roundf(v):
    """std::round(float): half away from zero."""
    t = np.trunc(v)
    return np.where(np.abs(v - t) >= np.float32(0.5), t + np.sign(v), t).astype(np.float32)


def grid_cells(pts_un, w, h, cam):
    """Frame::posInGrid / assignFeaturesToGrid, Core/Frame.cpp:75-89,231-241: cell x * 48 + y, -1 outside."""
    min_x, max_x, min_y, max_y = image_bounds(w, h, cam)
    w_inv = np.float32(FRAME_GRID_COLS) / np.float32(max_x - min_x)  # :65
    h_inv = np.float32(FRAME_GRID_ROWS) / np.float32(max_y - min_y)  # :66
    pts_un = np.asarray(pts_un, np.float32)
    gx = _roundf((pts_un[:, 0] - np.float32(min_x)) * w_inv).astype(np.int64)
    gy = _roundf((pts_un[:, 1] - np.float32(min_y)) * h_inv).astype(np.int64)
    ok = (gx >= 0) & (gx < FRAME_GRID_COLS) & (gy >= 0) & (gy < FRAME_GRID_ROWS)
    return np.where(ok, gx * FRAME_GRID_ROWS + gy, -1).astype(np.int32)


def unproject(pts_raw, pts_un, depth, cam):
    """Frame::uprojectCamera, Core/Frame.cpp:91-117: z = depth at the raw (truncated) position, x / y from the
    undistorted one, float arithmetic; (0, 0, 0) unless z > 0."""
    pts_raw = np.asarray(pts_raw, np.float32)
    pts_un = np.asarray(pts_un, np.float32)
    z = depth[pts_raw[:, 1].astype(np.int32), pts_raw[:, 0].astype(np.int32)].astype(np.float32)
    invfx = np.float32(1.0) / np.float32(cam["fx"])
    invfy = np.float32(1.0) / np.float32(cam["fy"])
    x = (pts_un[:, 0] - np.float32(cam["cx"])) * z * invfx
    y = (pts_un[:, 1] - np.float32(cam["cy"])) * z * invfy
    out = np.stack([x, y, z], 1).astype(np.float32)
    out[~(z > 0)] = 0
    return out
