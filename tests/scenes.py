"""Procedural inputs shared by the tests (no reference assets needed)."""
import numpy as np

import oracle_binding as ob


def random_rays(n, seed=0, scale=1.0):
    rng = np.random.default_rng(seed)
    rays = np.zeros((n, 8), np.float32)
    rays[:, :3] = rng.uniform(-0.9, 0.9, (n, 3)) * scale + [0, -0.5 * scale, 0]
    d = rng.normal(size=(n, 3))
    rays[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    return rays


def camera_rays(pc, W, H):
    """Primary rays of main() (raytrace.comp:541-556) in numpy float32 — close to, not bit-equal with, the kernels' rays;
    only used as a fixed ray batch."""
    f = np.frombuffer(bytes(pc), np.float32)
    rot = f[:16].reshape(4, 4).T[:3, :3]
    pos, near, aspect, fov = f[16:19], f[19], f[20], f[21]
    ph = np.float32(near * np.tan(np.radians(fov * 0.5)) * 2)
    pw = np.float32(ph * aspect)
    ys, xs = np.mgrid[0:H, 0:W].astype(np.float32)
    pt = np.stack([-pw / 2 + pw * xs / W, -ph / 2 + ph * ys / H, np.full_like(xs, 0.1)], -1)
    d = pt / np.linalg.norm(pt, axis=-1, keepdims=True)
    d = d @ rot.T
    rays = np.zeros((H * W, 8), np.float32)
    rays[:, :3] = pos
    rays[:, 4:7] = d.reshape(-1, 3)
    return rays


def soup_points(n, seed=0, edge=0.05):
    rng = np.random.default_rng(seed)
    c = rng.uniform(-1, 1, (n, 1, 3)).astype(np.float32)
    p = (c + rng.uniform(-edge, edge, (n, 3, 3)).astype(np.float32)).astype(np.float32)
    pts = np.zeros((n, 3, 8), np.float32)
    pts[:, :, :3] = p
    nrm = np.cross(p[:, 1] - p[:, 0], p[:, 2] - p[:, 0])
    nrm /= np.maximum(np.linalg.norm(nrm, axis=1, keepdims=True), 1e-20)
    pts[:, :, 4:7] = nrm[:, None, :]
    return pts


def soup_blobs(n, seed=0, edge=0.05):
    b = ob.OracleBuilder()
    b.push_default_materials()
    b.add_mesh(soup_points(n, seed, edge), material=0)
    return b.blobs()
