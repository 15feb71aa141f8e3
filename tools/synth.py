"""Synthetic RGB-D inputs shaped like the reference's datasets.

Test/bench infrastructure.  A textured box room is ray-cast per pixel (numpy),
giving a gray image (the reference converts BGR->gray before extraction,
Core/Frame.cpp:47) and a metric depth map (Core/Frame.cpp:48).  Intrinsics are
the ones hard-coded in the reference's IO classes:
  TUM fr1   IO/DatasetTUM.cpp:70-72      ICL-NUIM  IO/DatasetICL.cpp:37-39
  CORBS     IO/DatasetCORBS.cpp:37-39 (x3 for the Kinect-v2-shaped 1080p config)
Also: direct synthesis of descriptor/3D-point sets for the loop-verification
configs (4, 5) where rendering 4096 views is unnecessary.
"""
import numpy as np

CAMS = {
    "tum_fr1": dict(w=640, h=480, fx=517.306408, fy=516.469215, cx=318.643040, cy=255.313989, factor=5000.0),
    "icl": dict(w=640, h=480, fx=481.2, fy=-480.0, cx=319.5, cy=239.5, factor=5000.0),
    "kinect2": dict(w=1920, h=1080, fx=468.6 * 3, fy=468.61 * 3, cx=318.27 * 3, cy=243.99 * 3, factor=5000.0),
}


def make_texture(rng, size=1024):
    """Random rectangles / discs / checker patches, blurred, u8."""
    import cv2

    tex = np.full((size, size), 128, np.float32)
    for _ in range(900):
        kind = rng.integers(0, 3)
        x, y = rng.integers(0, size, 2)
        s = int(rng.integers(6, 70))
        val = float(rng.integers(10, 246))
        if kind == 0:
            tex[y : y + s, x : x + int(s * rng.uniform(0.4, 2.0))] = val
        elif kind == 1:
            cv2.circle(tex, (int(x), int(y)), s // 2, val, -1)
        else:
            c = int(rng.integers(3, 9))
            yy, xx = np.mgrid[0:s, 0:s]
            patch = np.where(((yy // c) + (xx // c)) % 2 == 0, val, 255.0 - val).astype(np.float32)
            sub = tex[y : y + s, x : x + s]
            sub[...] = patch[: sub.shape[0], : sub.shape[1]]
    tex = cv2.GaussianBlur(tex, (0, 0), 0.8)
    return np.clip(tex, 0, 255).astype(np.float32)


class Room:
    """Axis-aligned box room [-hx,hx] x [-hy,hy] x [-hz,hz], one texture per wall."""

    def __init__(self, seed=1234, half=(3.0, 2.0, 3.0), tex_size=1024):
        self.rng = np.random.default_rng(seed)
        self.half = np.array(half, np.float64)
        self.tex = [make_texture(self.rng, tex_size) for _ in range(6)]
        self.tex_size = tex_size

    def render(self, cam, R_wc, t_wc, noise_sigma=0.0, invalid_frac=0.0, max_depth=6.0, rng=None):
        """R_wc, t_wc: camera-to-world rotation / camera centre.  Returns gray u8, depth float32 (metres, 0 = invalid)."""
        w, h = cam["w"], cam["h"]
        u, v = np.meshgrid(np.arange(w, dtype=np.float64), np.arange(h, dtype=np.float64))
        d_c = np.stack([(u - cam["cx"]) / cam["fx"], (v - cam["cy"]) / cam["fy"], np.ones_like(u)], -1)
        d_w = d_c @ R_wc.T
        o = np.asarray(t_wc, np.float64)
        best_t = np.full((h, w), np.inf)
        gray = np.zeros((h, w), np.float32)
        face = 0
        for axis in range(3):
            for sign in (-1.0, 1.0):
                with np.errstate(divide="ignore", invalid="ignore"):
                    t = (sign * self.half[axis] - o[axis]) / d_w[..., axis]
                p = o + d_w * t[..., None]
                a, b = [k for k in range(3) if k != axis]
                inside = (t > 1e-6) & (np.abs(p[..., a]) <= self.half[a] + 1e-9) & (np.abs(p[..., b]) <= self.half[b] + 1e-9) & (t < best_t)
                if inside.any():
                    s = self.tex_size / (2.0 * max(self.half))  # texels per metre
                    tu = (p[..., a] + self.half[a]) * s
                    tv = (p[..., b] + self.half[b]) * s
                    tu = np.clip(tu, 0, self.tex_size - 1.001)
                    tv = np.clip(tv, 0, self.tex_size - 1.001)
                    iu, iv = tu.astype(np.int64), tv.astype(np.int64)
                    fu, fv = (tu - iu).astype(np.float32), (tv - iv).astype(np.float32)
                    T = self.tex[face]
                    val = (T[iv, iu] * (1 - fu) * (1 - fv) + T[iv, iu + 1] * fu * (1 - fv) + T[iv + 1, iu] * (1 - fu) * fv + T[iv + 1, iu + 1] * fu * fv)
                    gray = np.where(inside, val, gray)
                    best_t = np.where(inside, t, best_t)
                face += 1
        depth = np.where(np.isfinite(best_t), best_t, 0.0)  # z along camera axis since d_c.z == 1
        rng = rng or self.rng
        if noise_sigma > 0:
            gray = gray + rng.normal(0, noise_sigma, gray.shape).astype(np.float32)
        gray8 = np.clip(np.rint(gray), 0, 255).astype(np.uint8)
        depth = depth.astype(np.float32)
        depth[depth > max_depth] = 0.0
        if invalid_frac > 0:
            depth[rng.random(depth.shape) < invalid_frac] = 0.0
        # quantise like a u16 depth PNG / factor, then convertTo(CV_32F, 1/factor)
        d16 = np.clip(np.rint(depth * cam["factor"]), 0, 65535).astype(np.uint16)
        depth = d16.astype(np.float32) * np.float32(1.0 / cam["factor"])
        return gray8, depth


def rot_xyz(rx, ry, rz):
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def xyz_trajectory(n, seed=0, amp=0.2, rot_deg=5.0):
    """fr1/xyz-like: sinusoidal translation along x, then y, then z with small rotation jitter."""
    rng = np.random.default_rng(seed)
    poses = []
    jit = np.zeros(3)
    for i in range(n):
        ph = i / max(n, 1)
        t = np.zeros(3)
        seg = int(ph * 3) % 3
        t[seg] = amp * np.sin(2 * np.pi * (ph * 3 - seg) * 2)
        jit = 0.9 * jit + 0.1 * rng.normal(0, np.deg2rad(rot_deg), 3)
        poses.append((rot_xyz(*jit), t))
    return poses


def sequence(cam_name="tum_fr1", n=8, seed=1234, noise_sigma=2.0, invalid_frac=0.05):
    """Yield (gray, depth) frames of a synthetic sequence."""
    cam = CAMS[cam_name]
    # For ICL's negative fy the image is simply flipped vertically; keep it as is.
    room = Room(seed)
    rng = np.random.default_rng(seed + 1)
    frames = []
    for R, t in xyz_trajectory(n, seed + 2):
        frames.append(room.render(cam, R, t, noise_sigma, invalid_frac, rng=rng))
    return cam, frames


def unproject(cam, kps_xy, depth):
    """Frame::uprojectCamera (Core/Frame.cpp:91-117) without lens distortion:
    z = depth[int(v), int(u)]; x = (u-cx)*z*invfx; y = (v-cy)*z*invfy, float32; (0,0,0) if z <= 0."""
    u = kps_xy[:, 0].astype(np.float32)
    v = kps_xy[:, 1].astype(np.float32)
    z = depth[v.astype(np.int32), u.astype(np.int32)].astype(np.float32)
    invfx = np.float32(1.0) / np.float32(cam["fx"])
    invfy = np.float32(1.0) / np.float32(cam["fy"])
    x = (u - np.float32(cam["cx"])) * z * invfx
    y = (v - np.float32(cam["cy"])) * z * invfy
    out = np.stack([x, y, z], 1).astype(np.float32)
    out[~(z > 0)] = 0
    return out


def synth_pair_sets(n_kf, n_feat, seed=5678, corr_frac=0.4, flip_p=0.08, overlap_frac=1.0):
    """Config 4/5 direct synthesis: query descriptors/points + n_kf keyframes.
    A fraction of keyframes (overlap_frac) share true correspondences with the query."""
    rng = np.random.default_rng(seed)
    qd = rng.integers(0, 256, (n_feat, 32), dtype=np.uint8)
    z = rng.uniform(0.5, 4.5, n_feat)
    qp = np.stack([rng.uniform(-1, 1, n_feat) * z, rng.uniform(-0.8, 0.8, n_feat) * z, z], 1).astype(np.float32)
    kd = rng.integers(0, 256, (n_kf, n_feat, 32), dtype=np.uint8)
    zk = rng.uniform(0.5, 4.5, (n_kf, n_feat))
    kp = np.stack([rng.uniform(-1, 1, (n_kf, n_feat)) * zk, rng.uniform(-0.8, 0.8, (n_kf, n_feat)) * zk, zk], 2).astype(np.float32)
    T = np.tile(np.eye(4), (n_kf, 1, 1))
    nc = int(corr_frac * n_feat)
    for k in range(n_kf):
        if rng.random() > overlap_frac:
            continue
        ang = rng.uniform(-1, 1, 3) * np.deg2rad(15) / np.sqrt(3)
        R = rot_xyz(*ang)
        t = rng.uniform(-0.3, 0.3, 3) / np.sqrt(3)
        T[k, :3, :3], T[k, :3, 3] = R, t
        qi = rng.permutation(n_feat)[:nc]
        ki = rng.permutation(n_feat)[:nc]
        flips = rng.random((nc, 256)) < flip_p
        kd[k, ki] = qd[qi] ^ np.packbits(flips, axis=1)
        X = qp[qi].astype(np.float64)
        Y = X @ R.T + t
        Y += rng.normal(0, 1, Y.shape) * (0.0012 * Y[:, 2:3] ** 2)
        kp[k, ki] = Y.astype(np.float32)
    return qd, qp, kd, kp, T.astype(np.float32)
