"""tools/gen_tess_scene.py -- BASELINE.json configs[4]: the synthetic tessellated-mesh scene (BVH-traversal stress).

Seeded and reproducible; built directly as flat arrays (what tinyobj::LoadObj + Scene::load would hold) so that a
10 M-triangle instance never has to go through a ~1 GB OBJ text file:

    box   x in [-1,1], y in [0,1.6], z in [-1,1], open towards the camera (as the Cornell scenes), walls illum 7 (diffuse),
          floor illum 8 (Mixture: Kd .5, Ks .3, Ns 30)
    light one quad under the ceiling, Ke 35 (the only emitter)
    three displaced UV-spheres, radius 0.3, n x n tessellation (2 n (n-1) triangles each), smooth vertex normals,
          displacement = seeded value noise, amplitude 2 % of the radius:
          left  illum 3 (mirror) | middle illum 6 (glass, Ni 1.5, Tf 1) | right illum 2 (-> PhongBSDF: Ns 50, Kd .3, Ks .6)

    from gen_tess_scene import make_scene;  scene = make_scene(10_000_000, seed=1)     # -> trb.HostScene
    python tools/gen_tess_scene.py --tris 200000 --seed 1 --out /tmp/tess.npz [--obj /tmp/tess.obj]
"""
from __future__ import annotations

import argparse
import importlib
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

CAMERA = dict(eye=(0.0, 0.8, 3.8), at=(0.0, 0.8, 0.0), up=(0.0, 1.0, 0.0), fov=30.0)


def _value_noise(p, rng, cells=12):
    """trilinear value noise with a smoothstep fade on a seeded lattice; p: (N,3) in [-1,1] -> (N,) in [-1,1]"""
    lat = rng.random((cells + 1, cells + 1, cells + 1)).astype(np.float64) * 2 - 1
    q = (p * 0.5 + 0.5) * (cells - 1e-6)
    i = np.floor(q).astype(np.int64)
    f = q - i
    f = f * f * (3 - 2 * f)
    out = np.zeros(len(p))
    for dx in (0, 1):
        for dy in (0, 1):
            for dz in (0, 1):
                w = (f[:, 0] if dx else 1 - f[:, 0]) * (f[:, 1] if dy else 1 - f[:, 1]) * (f[:, 2] if dz else 1 - f[:, 2])
                out += w * lat[i[:, 0] + dx, i[:, 1] + dy, i[:, 2] + dz]
    return out


def _sphere(n, center, radius, rng, amp=0.02):
    """displaced UV sphere: poles are single vertices (fans), n-1 rings of n vertices between them"""
    th = np.pi * np.arange(1, n) / n                       # ring polar angles
    ph = 2 * np.pi * np.arange(n) / n
    T, P = np.meshgrid(th, ph, indexing="ij")
    unit = np.stack([np.sin(T) * np.cos(P), np.cos(T), np.sin(T) * np.sin(P)], -1).reshape(-1, 3)
    unit = np.concatenate([[[0.0, 1.0, 0.0]], unit, [[0.0, -1.0, 0.0]]])
    r = radius * (1 + amp * _value_noise(unit, rng))
    pos = center + unit * r[:, None]
    nr = n - 1                                             # rings
    idx = lambda i, j: 1 + i * n + (j % n)                 # noqa: E731
    jj = np.arange(n)
    tris = [np.stack([np.zeros(n, np.int64), idx(0, jj + 1), idx(0, jj)], 1)]            # top fan
    ii, jg = np.meshgrid(np.arange(nr - 1), jj, indexing="ij")
    a, b, c, d = idx(ii, jg), idx(ii, jg + 1), idx(ii + 1, jg), idx(ii + 1, jg + 1)
    tris.append(np.stack([a, b, d], -1).reshape(-1, 3))
    tris.append(np.stack([a, d, c], -1).reshape(-1, 3))
    last = 1 + nr * n
    tris.append(np.stack([np.full(n, last, np.int64), idx(nr - 1, jj), idx(nr - 1, jj + 1)], 1))  # bottom fan
    tris = np.concatenate(tris)
    # smooth normals: area-weighted average of the face normals around each vertex
    fn = np.cross(pos[tris[:, 1]] - pos[tris[:, 0]], pos[tris[:, 2]] - pos[tris[:, 0]])
    nrm = np.zeros_like(pos)
    for k in range(3):
        np.add.at(nrm, tris[:, k], fn)
    flip = np.sum(nrm * unit, 1) < 0                        # outward
    nrm[flip] *= -1
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    return pos.astype(np.float32), nrm.astype(np.float32), tris.astype(np.int32)


def _quad(p0, p1, p2, p3, normal):
    return (np.array([p0, p1, p2, p3], np.float32), np.tile(np.array(normal, np.float32), (4, 1)), np.array([[0, 1, 2], [0, 2, 3]], np.int32))


def make_scene(tris: int = 10_000_000, seed: int = 1):
    trb = importlib.import_module("bidirectional-path-tracing_b200")
    rng = np.random.default_rng(seed)
    n = max(8, int(np.sqrt(max(tris - 12, 0) / 6.0)))
    x0, x1, y0, y1, z0, z1 = -1.0, 1.0, 0.0, 1.6, -1.0, 1.0
    parts = [  # (name, material, geometry)
        ("light", 0, _quad((-0.25, 1.59, -0.25), (0.25, 1.59, -0.25), (0.25, 1.59, 0.25), (-0.25, 1.59, 0.25), (0, -1, 0))),
        ("backWall", 1, _quad((x0, y0, z0), (x1, y0, z0), (x1, y1, z0), (x0, y1, z0), (0, 0, 1))),
        ("ceiling", 1, _quad((x0, y1, z0), (x1, y1, z0), (x1, y1, z1), (x0, y1, z1), (0, -1, 0))),
        ("leftWall", 2, _quad((x0, y0, z0), (x0, y1, z0), (x0, y1, z1), (x0, y0, z1), (1, 0, 0))),
        ("rightWall", 3, _quad((x1, y0, z0), (x1, y0, z1), (x1, y1, z1), (x1, y1, z0), (-1, 0, 0))),
        ("floor", 4, _quad((x0, y0, z0), (x0, y0, z1), (x1, y0, z1), (x1, y0, z0), (0, 1, 0))),
        ("mirrorSphere", 5, _sphere(n, np.array([-0.55, 0.32, -0.3]), 0.3, rng)),
        ("glassSphere", 6, _sphere(n, np.array([0.0, 0.32, 0.35]), 0.3, rng)),
        ("phongSphere", 7, _sphere(n, np.array([0.55, 0.32, -0.3]), 0.3, rng)),
    ]
    mats = np.zeros(8, trb.MATERIAL_DTYPE)
    mats["ni"], mats["tf"] = 1.0, 1.0

    def setm(i, illum, kd=(0, 0, 0), ks=(0, 0, 0), ke=(0, 0, 0), ns=1.0, ni=1.0):
        mats[i]["illum"], mats[i]["kd"], mats[i]["ks"], mats[i]["ke"], mats[i]["ns"], mats[i]["ni"] = illum, kd, ks, ke, ns, ni
    setm(0, 7, kd=(0.78, 0.78, 0.78), ke=(35, 35, 35))
    setm(1, 7, kd=(0.73, 0.73, 0.73))
    setm(2, 7, kd=(0.63, 0.065, 0.05))
    setm(3, 7, kd=(0.14, 0.45, 0.091))
    setm(4, 8, kd=(0.5, 0.5, 0.5), ks=(0.3, 0.3, 0.3), ns=30.0)
    setm(5, 3, ks=(1, 1, 1))
    setm(6, 6, ks=(1, 1, 1), ni=1.5)
    setm(7, 2, kd=(0.3, 0.3, 0.3), ks=(0.6, 0.6, 0.6), ns=50.0)
    P, N, TV, TM, TS, names = [], [], [], [], [], []
    base = 0
    for s, (name, m, (pos, nrm, tri)) in enumerate(parts):
        P.append(pos); N.append(nrm); TV.append(tri + base)
        TM.append(np.full(len(tri), m, np.int32)); TS.append(np.full(len(tri), s, np.int32)); names.append(name)
        base += len(pos)
    tv = np.concatenate(TV)
    return trb.HostScene(np.concatenate(P), np.concatenate(N), tv, tv.copy(), np.concatenate(TM), np.concatenate(TS), mats,
                         n_shapes=len(parts), shape_names=names)


def make_sphere_light_scene(n: int = 20, seed: int = 3):
    """what DirectIntegrator (src/integrators/direct.h) is written for - SPHERE emitters, which no shipped scene
    has.  The same open box with two emissive UV spheres (exact, no displacement; different radii and radiances, so that
    selectEmitter and the per-emitter cone matter), a Phong sphere, a diffuse sphere and a glass sphere on a Mixture floor."""
    trb = importlib.import_module("bidirectional-path-tracing_b200")
    rng = np.random.default_rng(seed)
    x0, x1, y0, y1, z0, z1 = -1.0, 1.0, 0.0, 1.6, -1.0, 1.0
    parts = [
        ("lightA", 0, _sphere(n, np.array([-0.45, 1.25, -0.2]), 0.14, rng, amp=0.0)),
        ("lightB", 8, _sphere(n, np.array([0.5, 1.05, 0.25]), 0.09, rng, amp=0.0)),
        ("backWall", 1, _quad((x0, y0, z0), (x1, y0, z0), (x1, y1, z0), (x0, y1, z0), (0, 0, 1))),
        ("ceiling", 1, _quad((x0, y1, z0), (x1, y1, z0), (x1, y1, z1), (x0, y1, z1), (0, -1, 0))),
        ("leftWall", 2, _quad((x0, y0, z0), (x0, y1, z0), (x0, y1, z1), (x0, y0, z1), (1, 0, 0))),
        ("rightWall", 3, _quad((x1, y0, z0), (x1, y0, z1), (x1, y1, z1), (x1, y1, z0), (-1, 0, 0))),
        ("floor", 4, _quad((x0, y0, z0), (x0, y0, z1), (x1, y0, z1), (x1, y0, z0), (0, 1, 0))),
        ("phongSphere", 7, _sphere(n, np.array([-0.5, 0.3, -0.1]), 0.3, rng, amp=0.0)),
        ("diffuseSphere", 1, _sphere(n, np.array([0.45, 0.25, -0.35]), 0.25, rng, amp=0.0)),
        ("glassSphere", 6, _sphere(n, np.array([0.1, 0.2, 0.45]), 0.2, rng, amp=0.0)),
    ]
    mats = np.zeros(9, trb.MATERIAL_DTYPE)
    mats["ni"], mats["tf"] = 1.0, 1.0

    def setm(i, illum, kd=(0, 0, 0), ks=(0, 0, 0), ke=(0, 0, 0), ns=1.0, ni=1.0):
        mats[i]["illum"], mats[i]["kd"], mats[i]["ks"], mats[i]["ke"], mats[i]["ns"], mats[i]["ni"] = illum, kd, ks, ke, ns, ni
    setm(0, 7, kd=(0.78, 0.78, 0.78), ke=(30, 28, 24))
    setm(1, 7, kd=(0.73, 0.73, 0.73))
    setm(2, 7, kd=(0.63, 0.065, 0.05))
    setm(3, 7, kd=(0.14, 0.45, 0.091))
    setm(4, 8, kd=(0.5, 0.5, 0.5), ks=(0.3, 0.3, 0.3), ns=30.0)
    setm(5, 3, ks=(1, 1, 1))
    setm(6, 6, ks=(1, 1, 1), ni=1.5)
    setm(7, 2, kd=(0.3, 0.3, 0.3), ks=(0.6, 0.6, 0.6), ns=50.0)
    setm(8, 7, kd=(0.78, 0.78, 0.78), ke=(12, 20, 40))
    P, N, TV, TM, TS, names = [], [], [], [], [], []
    base = 0
    for s, (name, m, (pos, nrm, tri)) in enumerate(parts):
        P.append(pos); N.append(nrm); TV.append(tri + base)
        TM.append(np.full(len(tri), m, np.int32)); TS.append(np.full(len(tri), s, np.int32)); names.append(name)
        base += len(pos)
    tv = np.concatenate(TV)
    return trb.HostScene(np.concatenate(P), np.concatenate(N), tv, tv.copy(), np.concatenate(TM), np.concatenate(TS), mats,
                         n_shapes=len(parts), shape_names=names)


def _checker(w, h, a, b, cells, rng, byte_noise=12):
    """procedural 8-bit image -> the floats Tex::load would hold (pow(byte / 255, 2.2f), row 0 = bottom row of the file)"""
    yy, xx = np.mgrid[0:h, 0:w]
    pick = ((xx * cells // w + yy * cells // h) % 2)[..., None]
    img = np.where(pick == 0, np.array(a, np.float64), np.array(b, np.float64)) * (0.6 + 0.4 * xx[..., None] / max(w - 1, 1))
    byte = np.clip(np.round(img * 255) + rng.integers(-byte_noise, byte_noise + 1, img.shape), 0, 255).astype(np.uint8)
    import scene_to_obj
    return scene_to_obj.ppm_lut()[byte]


def make_textured_scene(n: int = 20, seed: int = 5):
    """bitmap textures (map_Kd / map_Ks -> BitmapTexture3f, core.h:525-585), which no shipped scene has.  The open box
    with a quad light; back wall Diffuse + map_Kd, floor Mixture + map_Kd + map_Ks, a Phong sphere + map_Kd (constant Ks), a
    Diffuse sphere + map_Kd, an untextured glass sphere; texture coordinates outside [0, 1) and negative ones exercise the wrap."""
    trb = importlib.import_module("bidirectional-path-tracing_b200")
    rng = np.random.default_rng(seed)
    x0, x1, y0, y1, z0, z1 = -1.0, 1.0, 0.0, 1.6, -1.0, 1.0
    parts = [  # (name, material, geometry, uv scale or None)
        ("light", 0, _quad((-0.25, 1.59, -0.25), (0.25, 1.59, -0.25), (0.25, 1.59, 0.25), (-0.25, 1.59, 0.25), (0, -1, 0)), None),
        ("backWall", 1, _quad((x0, y0, z0), (x1, y0, z0), (x1, y1, z0), (x0, y1, z0), (0, 0, 1)), 2.0),
        ("ceiling", 2, _quad((x0, y1, z0), (x1, y1, z0), (x1, y1, z1), (x0, y1, z1), (0, -1, 0)), None),
        ("leftWall", 3, _quad((x0, y0, z0), (x0, y1, z0), (x0, y1, z1), (x0, y0, z1), (1, 0, 0)), None),
        ("rightWall", 4, _quad((x1, y0, z0), (x1, y0, z1), (x1, y1, z1), (x1, y1, z0), (-1, 0, 0)), None),
        ("floor", 5, _quad((x0, y0, z0), (x0, y0, z1), (x1, y0, z1), (x1, y0, z0), (0, 1, 0)), 3.0),
        ("phongSphere", 6, _sphere(n, np.array([-0.5, 0.3, -0.1]), 0.3, rng, amp=0.0), 1.0),
        ("diffuseSphere", 7, _sphere(n, np.array([0.45, 0.25, -0.35]), 0.25, rng, amp=0.0), 1.0),
        ("glassSphere", 8, _sphere(n, np.array([0.1, 0.2, 0.45]), 0.2, rng, amp=0.0), None),
    ]
    centers = {"phongSphere": (-0.5, 0.3, -0.1), "diffuseSphere": (0.45, 0.25, -0.35)}
    mats = np.zeros(9, trb.MATERIAL_DTYPE)
    mats["ni"], mats["tf"] = 1.0, 1.0

    def setm(i, illum, kd=(0, 0, 0), ks=(0, 0, 0), ke=(0, 0, 0), ns=1.0, ni=1.0, kd_tex=-1, ks_tex=-1):
        mats[i]["illum"], mats[i]["kd"], mats[i]["ks"], mats[i]["ke"], mats[i]["ns"], mats[i]["ni"] = illum, kd, ks, ke, ns, ni
        mats[i]["tex"] = (kd_tex + 1) | ((ks_tex + 1) << 16)
    setm(0, 7, kd=(0.78, 0.78, 0.78), ke=(35, 35, 35))
    setm(1, 7, kd=(0.73, 0.73, 0.73), kd_tex=0)
    setm(2, 7, kd=(0.73, 0.73, 0.73))
    setm(3, 7, kd=(0.63, 0.065, 0.05))
    setm(4, 7, kd=(0.14, 0.45, 0.091))
    setm(5, 8, kd=(0.5, 0.5, 0.5), ks=(0.3, 0.3, 0.3), ns=30.0, kd_tex=1, ks_tex=2)
    setm(6, 2, kd=(0.3, 0.3, 0.3), ks=(0.6, 0.6, 0.6), ns=50.0, kd_tex=3)
    setm(7, 7, kd=(0.5, 0.5, 0.5), kd_tex=0)
    setm(8, 6, ks=(1, 1, 1), ni=1.5)
    textures = [_checker(64, 48, (0.9, 0.8, 0.3), (0.2, 0.3, 0.8), 8, rng), _checker(32, 32, (0.8, 0.8, 0.8), (0.25, 0.2, 0.15), 4, rng),
                _checker(16, 24, (0.7, 0.7, 0.7), (0.05, 0.05, 0.05), 2, rng), _checker(40, 20, (0.9, 0.2, 0.2), (0.9, 0.9, 0.9), 5, rng)]
    P, N, TV, TM, TS, TT, UV, names = [], [], [], [], [], [], [], []
    base, uv_base = 0, 0
    for s_, (name, m, (pos, nrm, tri), uvs) in enumerate(parts):
        P.append(pos); N.append(nrm); TV.append(tri + base)
        TM.append(np.full(len(tri), m, np.int32)); TS.append(np.full(len(tri), s_, np.int32)); names.append(name)
        if uvs is None:
            TT.append(np.full((len(tri), 3), -1, np.int32))
        else:
            if name in centers:
                d = pos.astype(np.float64) - np.array(centers[name])
                d /= np.linalg.norm(d, axis=1, keepdims=True)
                uv = np.stack([np.arctan2(d[:, 2], d[:, 0]) / (2 * np.pi), np.arccos(np.clip(d[:, 1], -1, 1)) / np.pi], 1)  # u in [-0.5, 0.5]
            else:
                uv = np.array([[0, 0], [1, 0], [1, 1], [0, 1]], np.float64) * uvs - 0.75
            UV.append(uv.astype(np.float32)); TT.append(tri + uv_base)
            uv_base += len(uv)
        base += len(pos)
    tv = np.concatenate(TV)
    return trb.HostScene(np.concatenate(P), np.concatenate(N), tv, tv.copy(), np.concatenate(TM), np.concatenate(TS), mats,
                         n_shapes=len(parts), shape_names=names, texcoords=np.concatenate(UV), tri_t=np.concatenate(TT), textures=textures)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--tris", type=int, default=200_000)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--out", default="")
    ap.add_argument("--obj", default="")
    a = ap.parse_args()
    sc = make_scene(a.tris, a.seed)
    print(f"{sc.n_tris} triangles, {len(sc.positions)} vertices, {sc.n_shapes} shapes")
    if a.out:
        sc.save_npz(a.out)
    if a.obj:
        import scene_to_obj
        scene_to_obj.write_obj(sc, a.obj)
