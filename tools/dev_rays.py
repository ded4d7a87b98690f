"""tools/dev_rays.py -- seeded ray sets used by the parity tests and golden generation:
camera rays, surface-origin rays in random directions, and visibility segments between surface points
(area-weighted, so many segments lie IN the plane of a wall: the numerically nasty case)."""
import numpy as np

FLT_MAX = np.finfo(np.float32).max


def surface_points(scene, n, rng):
    P = scene.positions
    a, b, c = P[scene.tri_v[:, 0]], P[scene.tri_v[:, 1]], P[scene.tri_v[:, 2]]
    area = 0.5 * np.linalg.norm(np.cross(b - a, c - a), axis=1).astype(np.float64)
    tri = rng.choice(len(area), size=n, p=area / area.sum())
    u = np.sqrt(rng.random(n)).astype(np.float32)
    v = rng.random(n).astype(np.float32)
    w0, w1, w2 = (1 - u), u * (1 - v), u * v
    p = (a[tri] * w0[:, None] + b[tri] * w1[:, None] + c[tri] * w2[:, None]).astype(np.float32)
    return p, tri


def camera_rays(cam_mats, eye, fov, w, h):
    """pixel-centre rays exactly as renderer.cpp:165-177 (float32 arithmetic)."""
    f = np.float32
    c2w = cam_mats[1]  # column-major: c2w[col][row]
    angle = np.tan(f(np.float32(3.14159265358979323846) / f(180.0) * f(fov) * f(0.5)), dtype=np.float32)
    aspect = f(w) / f(h)
    j, i = np.meshgrid(np.arange(w), np.arange(h))
    y = (f(1) - (i.astype(f) + f(0.5)) * (f(1) / f(h))) * f(2) - f(1)
    x = ((j.astype(f) + f(0.5)) * (f(1) / f(w))) * f(2) - f(1)
    px, py, pz = (x * angle * aspect).ravel(), (y * angle).ravel(), np.full(w * h, -1, f)
    d = np.stack([(c2w[0][k] * px + c2w[1][k] * py) + (c2w[2][k] * pz + c2w[3][k] * f(0)) for k in range(3)], 1).astype(f)
    inv = f(1) / np.sqrt((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2], dtype=f)
    d = d * inv[:, None]
    r = np.zeros((w * h, 8), f)
    r[:, 0:3] = np.asarray(eye, f); r[:, 3:6] = d; r[:, 6] = 1; r[:, 7] = 1000
    return r


def random_dir_rays(scene, n, rng):
    p, _ = surface_points(scene, n, rng)
    z = rng.random(n) * 2 - 1
    ph = rng.random(n) * 2 * np.pi
    s = np.sqrt(1 - z * z)
    r = np.zeros((n, 8), np.float32)
    r[:, 0:3] = p; r[:, 3:6] = np.stack([s * np.cos(ph), s * np.sin(ph), z], 1).astype(np.float32)
    r[:, 6] = 1e-8; r[:, 7] = FLT_MAX
    return r


def segment_rays(scene, n, rng):
    """visibilityQuery(start,end) rays (bdpt.h:498-505) in float32: d = (end-start)/len, [1e-8, len-1e-5]"""
    f = np.float32
    a, _ = surface_points(scene, n, rng)
    b, _ = surface_points(scene, n, rng)
    d = (b - a).astype(f)
    dist = np.sqrt((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2], dtype=f)
    ok = dist > 0
    a, b, d, dist = a[ok], b[ok], d[ok], dist[ok]
    d = (d / dist[:, None]).astype(f)
    r = np.zeros((len(a), 8), f)
    r[:, 0:3] = a; r[:, 3:6] = d; r[:, 6] = f(1e-8); r[:, 7] = dist - f(0.00001)
    return r, np.concatenate([a, b], 1)


def axis_rays(scene, n, rng):
    """rays with zero direction components (division by zero in the reference's slab test)"""
    p, _ = surface_points(scene, n, rng)
    r = np.zeros((n, 8), np.float32)
    r[:, 0:3] = p + (rng.random((n, 3)).astype(np.float32) - 0.5) * 0.2
    ax = rng.integers(0, 3, n); sg = rng.choice([-1.0, 1.0], n)
    r[np.arange(n), 3 + ax] = sg
    r[:, 6] = 1e-8; r[:, 7] = FLT_MAX
    return r
