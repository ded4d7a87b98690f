"""oracle/refh.py -- TEST INFRASTRUCTURE ONLY.

ctypes wrapper over oracle/_ref/libtr_ref.so, i.e. the UNMODIFIED reference compiled by oracle/Makefile
plus the batch harness oracle/ref_harness.cpp.  Only tests/, tools/gen_golden.py, __graft_entry__.smoke()
and bench.py's CPU-baseline leg may import this module; the product never does.
"""
from __future__ import annotations

import ctypes as C
import os
import tempfile
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
REF_LIB = HERE / "_ref" / "libtr_ref.so"
REF_BIN = HERE / "_ref" / "tinyrender_ref"
# the same sources with the reference's own Russian-roulette code compiled in (`make -C oracle ref_rr`: NO_RR 0, bdpt.h:18)
REF_RR_LIB = HERE / "_ref" / "libtr_ref_rr.so"


# ... and with one of its estimator A/B switches flipped (`make -C oracle ref_lt ref_pt`: LIGHT_TRACING 1 / PATH_TRACING 1, bdpt.h:16-17)
VARIANT_LIBS = {False: REF_LIB, True: REF_RR_LIB, "lt": HERE / "_ref" / "libtr_ref_lt.so", "pt": HERE / "_ref" / "libtr_ref_pt.so"}


def available(rr=False) -> bool:
    """rr: False = the reference as shipped, True = NO_RR 0, "lt" / "pt" = LIGHT_TRACING 1 / PATH_TRACING 1"""
    return VARIANT_LIBS[rr].exists()


def write_toml(path, objfile, width=800, height=600, spp=1, rr_depth=5, eye=(0.0, 0.8, 3.8), at=(0.0, 0.8, 0.0),
               up=(0.0, 1.0, 0.0), fov=30.0, type_="bdpt", rr_prob=0.95, extra="") -> str:
    """extra: further `[renderer]` lines, e.g. 'isExplicit = true\nmaxDepth = -1\n' for type_="path" (main.cpp:94-102)"""
    def v(x):
        return "[ " + ", ".join(repr(float(c)) for c in x) + " ]"
    txt = (f'[input]\nobjfile = "{objfile}"\n\n[camera]\neye = {v(eye)}\nat = {v(at)}\nup = {v(up)}\nfov = {float(fov)!r}\n\n'
           f'[film]\nwidth = {int(width)}\nheight = {int(height)}\n\n[renderer]\nrealtime = false\ntype = "{type_}"\n'
           f'rrDepth = {int(rr_depth)}\nrrProb = {float(rr_prob)!r}\nspp = {int(spp)}\n{extra}')
    Path(path).write_text(txt)
    return str(path)


_libs = {}


def _L(rr=False):
    if rr not in _libs:
        path = VARIANT_LIBS[rr]
        if not path.exists():
            raise FileNotFoundError(f"{path} missing: run `make -C oracle ref ref_rr ref_lt ref_pt` where /root/reference exists")
        L = C.CDLL(str(path))
        L.refh_open.restype = C.c_void_p
        L.refh_open.argtypes = [C.c_char_p]
        L.refh_render.restype = C.c_double
        L.refh_bvh_nodes.restype = C.c_int
        _libs[rr] = L
    return _libs[rr]


def _p(a):
    return C.c_void_p(a.ctypes.data) if a is not None else None


class RefHarness:
    """The reference renderer opened on a TOML scene file (stdout chatter of the reference is kept)."""

    def __init__(self, toml_path, rr=False):
        self.L = _L(rr)
        self.h = C.c_void_p(self.L.refh_open(str(toml_path).encode()))
        if not self.h:
            raise RuntimeError(f"reference failed to open {toml_path}")
        c = np.zeros(10, np.int64)
        self.L.refh_counts(self.h, _p(c))
        (self.n_shapes, self.n_materials, self.n_vfloats, self.n_nfloats, self.n_tris, self.n_emitters, self.width,
         self.height, self.spp, self.rr_depth) = [int(x) for x in c]

    @classmethod
    def open_scene(cls, objfile, rr=False, **kw):
        d = tempfile.mkdtemp(prefix="refh_")
        return cls(write_toml(os.path.join(d, "scene.toml"), objfile, **kw), rr=rr)

    def close(self):
        if self.h:
            self.L.refh_close(self.h)
            self.h = C.c_void_p()

    def scene_dump(self) -> dict:
        v = np.zeros(self.n_vfloats, np.float32)
        n = np.zeros(self.n_nfloats, np.float32)
        tv = np.zeros((self.n_tris, 3), np.int32)
        tn = np.zeros((self.n_tris, 3), np.int32)
        tm = np.zeros(self.n_tris, np.int32)
        ts = np.zeros(self.n_tris, np.int32)
        self.L.refh_scene_dump(self.h, _p(v), _p(n), _p(tv), _p(tn), _p(tm), _p(ts))
        m = np.zeros((self.n_materials, 16), np.float32)
        self.L.refh_materials(self.h, _p(m))
        return dict(positions=v.reshape(-1, 3), normals=n.reshape(-1, 3), tri_v=tv, tri_n=tn, tri_mat=tm, tri_shape=ts, materials=m)

    def emitters(self) -> dict:
        e = np.zeros((self.n_emitters, 5), np.float32)
        sizes = np.zeros(self.n_emitters, np.int32)
        self.L.refh_emitters(self.h, _p(e), _p(sizes))
        cdfs = []
        for i in range(self.n_emitters):
            c = np.zeros(int(sizes[i]), np.float32)
            self.L.refh_emitter_cdf(self.h, i, _p(c))
            cdfs.append(c)
        return dict(shape=e[:, 0].astype(np.int32), area=e[:, 1].copy(), radiance=e[:, 2:5].copy(), cdf=cdfs)

    def bvh_order(self) -> np.ndarray:
        o = np.zeros(self.n_tris, np.int32)
        self.L.refh_bvh_order(self.h, _p(o))
        return o

    def bvh_nodes(self):
        n = self.L.refh_bvh_nodes(self.h, None, None, 0)
        bbox = np.zeros((n, 6), np.float32)
        meta = np.zeros((n, 3), np.uint32)
        self.L.refh_bvh_nodes(self.h, _p(bbox), _p(meta), n)
        return bbox, meta

    def camera_matrices(self) -> np.ndarray:
        out = np.zeros((4, 4, 4), np.float32)
        self.L.refh_camera_matrices(self.h, _p(out))
        return out

    def intersect(self, rays8: np.ndarray):
        rays8 = np.ascontiguousarray(rays8, np.float32).reshape(-1, 8)
        n = len(rays8)
        ids = np.zeros((n, 3), np.int32)
        tuv = np.zeros((n, 3), np.float32)
        geo = np.zeros((n, 12), np.float32)
        self.L.refh_intersect(self.h, _p(rays8), C.c_int64(n), _p(ids), _p(tuv), _p(geo))
        return ids, tuv, geo

    def occluded(self, rays8: np.ndarray) -> np.ndarray:
        rays8 = np.ascontiguousarray(rays8, np.float32).reshape(-1, 8)
        out = np.zeros(len(rays8), np.uint8)
        self.L.refh_occluded(self.h, _p(rays8), C.c_int64(len(rays8)), _p(out))
        return out

    def visibility(self, segs6: np.ndarray) -> np.ndarray:
        segs6 = np.ascontiguousarray(segs6, np.float32).reshape(-1, 6)
        out = np.zeros(len(segs6), np.uint8)
        self.L.refh_visibility(self.h, _p(segs6), C.c_int64(len(segs6)), _p(out))
        return out

    def bsdf(self, mat: int, q8: np.ndarray) -> np.ndarray:
        q8 = np.ascontiguousarray(q8, np.float32).reshape(-1, 8)
        out = np.zeros((len(q8), 12), np.float32)
        self.L.refh_bsdf(self.h, int(mat), _p(q8), C.c_int64(len(q8)), _p(out))
        return out

    def bdpt_paths(self, rays8: np.ndarray, seeds: np.ndarray, max_splats: int):
        rays8 = np.ascontiguousarray(rays8, np.float32).reshape(-1, 8)
        seeds = np.ascontiguousarray(seeds, np.int32)
        n = len(rays8)
        li = np.zeros((n, 3), np.float32)
        pix = np.zeros((n, max_splats), np.int32)
        rgb = np.zeros((n, max_splats, 3), np.float32)
        nd = np.zeros(n, np.int32)
        self.L.refh_bdpt_paths(self.h, _p(rays8), _p(seeds), C.c_int64(n), int(max_splats), _p(li), _p(pix), _p(rgb), _p(nd))
        return li, pix, rgb, nd

    def integrator_paths(self, rays8: np.ndarray, seeds: np.ndarray):
        """Integrator::render of whatever integrator the TOML named (used for type = "path")"""
        rays8 = np.ascontiguousarray(rays8, np.float32).reshape(-1, 8)
        seeds = np.ascontiguousarray(seeds, np.int32)
        n = len(rays8)
        li = np.zeros((n, 3), np.float32)
        nd = np.zeros(n, np.int32)
        self.L.refh_integrator_paths(self.h, _p(rays8), _p(seeds), C.c_int64(n), _p(li), _p(nd))
        return li, nd

    def shape_spheres(self, n_shapes: int) -> np.ndarray:
        """Scene::getShapeCenter / getShapeRadius per shape: (n_shapes, 4) = centre xyz, radius"""
        out = np.zeros((n_shapes, 4), np.float32)
        self.L.refh_shape_spheres(self.h, _p(out))
        return out

    def render(self):
        img = np.zeros((self.height, self.width, 3), np.float32)
        sec = self.L.refh_render(self.h, _p(img))
        return img, float(sec)


def sampler_stream(seed: int, n: int) -> np.ndarray:
    out = np.zeros(n, np.float32)
    _L().refh_sampler_stream(int(seed), C.c_int64(n), _p(out))
    return out
