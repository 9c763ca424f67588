"""Synthetic scenes and cameras for parity tests and bench.py (numpy only; SURVEY 8d / BASELINE.md section 4).

Camera construction follows the reference trainer (cuda/trainer.cu:1299-1339): proj from the field of view with
znear 0.01 / zfar 100, view = [R|t] row-major.  Deviation from the survey's spec, stated here: the RNG is numpy's
PCG64 seeded with 0x3D65 + config index (one generator per scene) instead of xoshiro256**; distributions are as
specified.
"""
import numpy as np

CONFIGS = {
    # name: (N, width, height, l_max)
    "C1": (1000, 100, 80, 0),
    "C2": (1_000_000, 1297, 840, 3),
    "C3": (3_000_000, 1237, 822, 3),
    "C4": (6_000_000, 1920, 1080, 3),
    "C5": (10_000_000, 3840, 2160, 3),
}
NEAR_THRESH, MH_DIST, CULL_PADDING = 0.3, 3.0, 100   # config/base.yaml:9-11


def projection_matrix(fx, fy, width, height, znear=0.01, zfar=100.0):
    """cuda/trainer.cu:1299-1320"""
    tan_x = np.tan(np.arctan(width / (2 * fx)))
    tan_y = np.tan(np.arctan(height / (2 * fy)))
    top, right = tan_y * znear, tan_x * znear
    p = np.zeros(16, np.float32)
    p[0] = 2 * znear / (2 * right)
    p[5] = 2 * znear / (2 * top)
    p[14] = 1.0
    p[10] = zfar / (zfar - znear)
    p[11] = -(zfar * znear) / (zfar - znear)
    return p.reshape(4, 4)


def look_at(cam_pos, target=(0, 0, 0), up=(0, 1, 0)):
    c = np.asarray(cam_pos, np.float64)
    f = np.asarray(target, np.float64) - c
    f /= np.linalg.norm(f)
    r = np.cross(np.asarray(up, np.float64), f)
    r /= np.linalg.norm(r)
    u = np.cross(f, r)
    R = np.stack([r, u, f])
    V = np.eye(4)
    V[:3, :3] = R
    V[:3, 3] = -R @ c
    return V.astype(np.float32), c.astype(np.float32)


def orbit_camera(width, height, azimuth_deg=0.0, radius=4.0, focal_scale=1.2):
    """PINHOLE, fx = fy = 1.2 W, on the radius-4 orbit; azimuth 0 = (0,0,-4) looking at the origin."""
    a = np.deg2rad(azimuth_deg)
    pos = (radius * np.sin(a), 0.0, -radius * np.cos(a))
    view, campos = look_at(pos)
    fx = fy = focal_scale * width
    return dict(view=view, proj=projection_matrix(fx, fy, width, height), campos=campos, fx=float(fx), fy=float(fy),
                width=width, height=height)


def synthetic_gaussians(N, l_max=3, seed=0x3D65, sh_alloc_lmax=None):
    """70 % uniform in the ball of radius 2.5, 30 % in 4096 clusters (sigma 0.05); log-scale ~ N(log s0, 0.6^2),
    s0 = 0.9 (N/1e6)^(-1/3) 0.012; unit quaternions; opacity logit ~ N(0, 2^2); band0 ~ U(-1.5, 1.5);
    higher SH ~ N(0, 0.15^2)."""
    rng = np.random.default_rng(seed)
    n_u = int(0.7 * N)
    d = rng.standard_normal((n_u, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True) + 1e-12
    rad = 2.5 * rng.random(n_u).astype(np.float32) ** (1.0 / 3.0)
    xyz_u = d * rad[:, None]
    n_c = N - n_u
    nclus = 4096
    cd = rng.standard_normal((nclus, 3)).astype(np.float32)
    cd /= np.linalg.norm(cd, axis=1, keepdims=True) + 1e-12
    centers = cd * (2.3 * rng.random(nclus).astype(np.float32) ** (1.0 / 3.0))[:, None]
    which = rng.integers(0, nclus, n_c)
    xyz_c = centers[which] + 0.05 * rng.standard_normal((n_c, 3)).astype(np.float32)
    xyz = np.concatenate([xyz_u, xyz_c]).astype(np.float32)
    perm = rng.permutation(N)
    xyz = xyz[perm]
    s0 = 0.9 * (max(N, 1) / 1e6) ** (-1.0 / 3.0) * 0.012
    scale = (np.log(s0) + 0.6 * rng.standard_normal((N, 3))).astype(np.float32)
    quat = rng.standard_normal((N, 4)).astype(np.float32)
    quat /= np.linalg.norm(quat, axis=1, keepdims=True) + 1e-12
    opacity = (2.0 * rng.standard_normal(N)).astype(np.float32)
    band0 = rng.uniform(-1.5, 1.5, (N, 3)).astype(np.float32)
    la = l_max if sh_alloc_lmax is None else sh_alloc_lmax
    nsh = (la + 1) ** 2 - 1
    sh = (0.15 * rng.standard_normal((N, nsh, 3))).astype(np.float32) if nsh > 0 else None
    return dict(xyz=xyz, scale=scale, quaternion=quat, opacity=opacity, band0=band0, sh=sh)


def frustum_gaussians(N, cam, depth_range=(8.0, 14.0), l_max=0, seed=0x3D65, scale_mean=np.log(0.08)):
    """C1 companion: N Gaussians seeded inside the camera frustum at the given depth range."""
    rng = np.random.default_rng(seed)
    W, H, fx, fy = cam["width"], cam["height"], cam["fx"], cam["fy"]
    z = rng.uniform(depth_range[0], depth_range[1], N)
    u = rng.uniform(0, W, N)
    v = rng.uniform(0, H, N)
    xc = (u - W / 2) * z / fx
    yc = (v - H / 2) * z / fy
    pc = np.stack([xc, yc, z], 1)
    V = cam["view"].astype(np.float64)
    R, t = V[:3, :3], V[:3, 3]
    xyz = ((pc - t) @ R).astype(np.float32)   # R^T (pc - t)
    scale = (scale_mean + 0.5 * rng.standard_normal((N, 3))).astype(np.float32)
    quat = rng.standard_normal((N, 4)).astype(np.float32)
    quat /= np.linalg.norm(quat, axis=1, keepdims=True)
    opacity = (1.5 * rng.standard_normal(N)).astype(np.float32)
    band0 = rng.uniform(-1.5, 1.5, (N, 3)).astype(np.float32)
    nsh = (l_max + 1) ** 2 - 1
    sh = (0.15 * rng.standard_normal((N, nsh, 3))).astype(np.float32) if nsh > 0 else None
    return dict(xyz=xyz, scale=scale, quaternion=quat, opacity=opacity, band0=band0, sh=sh)


def grad_image(width, height, seed=7):
    """Stand-in for the loss gradient: N(0,1)/(3 W H)  (SURVEY 8d)."""
    rng = np.random.default_rng(seed)
    return (rng.standard_normal((height, width, 3)) / (3.0 * width * height)).astype(np.float32)


# ---- the reference's COLMAP fixture, decoded (/root/reference/test_data/*.bin, 260 bytes) ----
def qvec_to_rotmat(q):
    w, x, y, z = q
    return np.array([[1 - 2 * y * y - 2 * z * z, 2 * x * y - 2 * w * z, 2 * z * x + 2 * w * y],
                     [2 * x * y + 2 * w * z, 1 - 2 * x * x - 2 * z * z, 2 * y * z - 2 * w * x],
                     [2 * z * x - 2 * w * y, 2 * y * z + 2 * w * x, 1 - 2 * x * x - 2 * y * y]], np.float64)


def test_data_camera():
    """cameras.bin: SIMPLE_PINHOLE 100x80 (f 150.5, cx 50.2, cy 40.8); images.bin: q=(0.8,0.1,0.2,0.3) (non-unit),
    t=(5.1,6.2,7.3).  fx, fy = params[0], params[1] regardless of model (cuda/raster.cu:89-90, Q16) -> fy = 50.2."""
    W, H = 100, 80
    fx, fy = 150.5, 50.2
    R = qvec_to_rotmat((0.8, 0.1, 0.2, 0.3))   # src/colmap.cpp QvecToRotMat does not normalise
    t = np.array([5.1, 6.2, 7.3])
    V = np.eye(4)
    V[:3, :3] = R
    V[:3, 3] = t
    campos = -R.T @ t                           # Image::CamPos
    return dict(view=V.astype(np.float32), proj=projection_matrix(fx, fy, W, H), campos=campos.astype(np.float32),
                fx=fx, fy=fy, width=W, height=H)


def test_data_gaussian():
    """points3D.bin: one point (1.1, 2.2, 3.3) rgb (10,20,30) -> one Gaussian (deg 0)."""
    xyz = np.array([[1.1, 2.2, 3.3]], np.float32)
    rgb = np.array([[10, 20, 30]], np.float32) / 255.0
    band0 = (rgb - 0.5) / 0.28209479177387814
    return dict(xyz=xyz, scale=np.log(np.full((1, 3), 0.05, np.float32)), quaternion=np.array([[1, 0, 0, 0]], np.float32),
                opacity=np.array([-1.3862944], np.float32), band0=band0.astype(np.float32), sh=None)
