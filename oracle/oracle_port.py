"""oracle/oracle_port.py -- TEST INFRASTRUCTURE ONLY: ctypes wrapper over oracle/_build/libtr_oracle.so, the plain
C++ restatement of the reference's BDPT path (oracle/bdpt_oracle.cpp).  Builds the library on first use (gcc only).
Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline leg may import this; the product never does."""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB = HERE / "_build" / "libtr_oracle.so"
_lib = None


def lib():
    global _lib
    if _lib is None:
        src = HERE / "bdpt_oracle.cpp"
        if not LIB.exists() or LIB.stat().st_mtime < src.stat().st_mtime:
            subprocess.run(["make", "-C", str(HERE), "port"], check=True, stdout=subprocess.DEVNULL)
        L = C.CDLL(str(LIB))
        L.orc_create.restype = C.c_void_p
        L.orc_render.restype = C.c_double
        _lib = L
    return _lib


def _p(a):
    return C.c_void_p(a.ctypes.data) if a is not None else None


class Oracle:
    """scene: any object with positions, normals, tri_v, tri_n, tri_mat, tri_shape, materials (HostScene layout)."""

    def __init__(self, scene):
        L = lib()
        m = scene.materials
        mats = np.zeros((len(m), 16), np.float32)
        mats[:, 0:3], mats[:, 3:6], mats[:, 6:9], mats[:, 9:12] = m["kd"], m["ks"], m["ke"], m["tf"]
        mats[:, 12], mats[:, 13], mats[:, 14] = m["ns"], m["ni"], m["illum"]
        textures = list(getattr(scene, "textures", []))
        if textures:
            mats[:, 15] = m["tex"]
        self.n_tris, self.n_mats = len(scene.tri_mat), len(m)
        self._keep = [np.ascontiguousarray(x) for x in (scene.positions, scene.normals, scene.tri_v, scene.tri_n, scene.tri_mat, scene.tri_shape, mats)]
        k = self._keep
        self.L = L
        tc = np.ascontiguousarray(getattr(scene, "texcoords", np.zeros((0, 2), np.float32)), np.float32)
        tt = np.ascontiguousarray(getattr(scene, "tri_t", np.zeros((0, 3), np.int32)), np.int32)
        textures = [np.ascontiguousarray(t, np.float32) for t in textures]
        ptrs = (C.c_void_p * max(len(textures), 1))(*[t.ctypes.data for t in textures])
        wh = np.array([[t.shape[1], t.shape[0]] for t in textures], np.int32).reshape(-1, 2)
        self._keep += [tc, tt, textures, ptrs, wh]
        self.h = C.c_void_p(L.orc_create(_p(k[0]), C.c_size_t(len(k[0])), _p(k[1]), C.c_size_t(len(k[1])), _p(k[2]), _p(k[3]), _p(k[4]), _p(k[5]),
                                         C.c_size_t(self.n_tris), _p(k[6]), C.c_size_t(self.n_mats),
                                         _p(tc), C.c_size_t(len(tc)), _p(tt), ptrs, _p(wh), C.c_size_t(len(textures))))
        self.width, self.height = 800, 600

    def close(self):
        if self.h:
            self.L.orc_destroy(self.h)
            self.h = C.c_void_p()

    def set_config(self, desc):
        """desc: the product's RenderDesc (eye/at/up/fov/width/height/spp/rr_depth)"""
        eye, at, up = (np.array(list(x), np.float32) for x in (desc.eye, desc.at, desc.up))
        self.L.orc_set_config(self.h, _p(eye), _p(at), _p(up), C.c_float(desc.fov_deg), int(desc.width), int(desc.height), int(desc.spp), int(desc.rr_depth))
        self.width, self.height = int(desc.width), int(desc.height)

    def set_rr(self, on: bool, max_depth: int = 0):
        """the reference's Russian-roulette code (NO_RR 0); max_depth > 0 adds the device's hard cap, 0 = none as in the reference"""
        self.L.orc_set_rr(self.h, int(bool(on)), int(max_depth))

    def set_path(self, is_explicit=True, max_depth=-1, rr_prob=0.95, emitter_samples=1, bsdf_samples=0, depth_cap=0):
        """PathTracerIntegrator settings (main.cpp:94-102); depth_cap > 0 = the device's cap on the recursion, 0 = none (reference)"""
        self.L.orc_set_path(self.h, int(bool(is_explicit)), int(max_depth), C.c_float(rr_prob), int(emitter_samples), int(bsdf_samples), int(depth_cap))

    def path_paths(self, rays8, seeds=None, injected=None):
        rays8 = np.ascontiguousarray(rays8, np.float32).reshape(-1, 8)
        n = len(rays8)
        li, nd = np.zeros((n, 3), np.float32), np.zeros(n, np.int32)
        if seeds is not None:
            seeds = np.ascontiguousarray(seeds, np.int32)
            self.L.orc_path_paths(self.h, _p(rays8), _p(seeds), None, C.c_int64(0), C.c_int64(n), _p(li), _p(nd))
        else:
            injected = np.ascontiguousarray(injected, np.float32)
            self.L.orc_path_paths(self.h, _p(rays8), None, _p(injected), C.c_int64(injected.shape[1]), C.c_int64(n), _p(li), _p(nd))
        return li, nd

    STRATEGIES = {"mis": 0, "area": 1, "solidAngle": 2, "cosineHemisphere": 3, "bsdf": 4}

    def set_direct(self, strategy="mis", emitter_samples=1, bsdf_samples=1):
        """DirectIntegrator settings (main.cpp:88-93)"""
        self.L.orc_set_direct(self.h, int(self.STRATEGIES.get(strategy, strategy)), int(emitter_samples), int(bsdf_samples))

    def direct_paths(self, rays8, seeds=None, injected=None):
        rays8 = np.ascontiguousarray(rays8, np.float32).reshape(-1, 8)
        n = len(rays8)
        li, nd = np.zeros((n, 3), np.float32), np.zeros(n, np.int32)
        if seeds is not None:
            seeds = np.ascontiguousarray(seeds, np.int32)
            self.L.orc_direct_paths(self.h, _p(rays8), _p(seeds), None, C.c_int64(0), C.c_int64(n), _p(li), _p(nd))
        else:
            injected = np.ascontiguousarray(injected, np.float32)
            self.L.orc_direct_paths(self.h, _p(rays8), None, _p(injected), C.c_int64(injected.shape[1]), C.c_int64(n), _p(li), _p(nd))
        return li, nd

    def shape_spheres(self, n_shapes):
        out = np.zeros((n_shapes, 4), np.float32)
        self.L.orc_shape_spheres(self.h, _p(out))
        return out

    def bvh_order(self):
        o = np.zeros(self.n_tris, np.int32)
        self.L.orc_bvh_order(self.h, _p(o))
        return o

    def emitters(self):
        cap = 64
        shape, area, rad = np.zeros(cap, np.int32), np.zeros(cap, np.float32), np.zeros((cap, 3), np.float32)
        cdf = np.zeros(self.n_tris + 2 * cap, np.float32)
        n = self.L.orc_emitters(self.h, _p(shape), _p(area), _p(rad), _p(cdf), cap)
        return dict(n=n, shape=shape[:n], area=area[:n], radiance=rad[:n], cdf=cdf)

    def camera_matrices(self):
        out = np.zeros((4, 4, 4), np.float32)
        self.L.orc_camera_matrices(self.h, _p(out))
        return out

    def intersect(self, rays8):
        rays8 = np.ascontiguousarray(rays8, np.float32).reshape(-1, 8)
        ids, tuv = np.zeros((len(rays8), 3), np.int32), np.zeros((len(rays8), 3), np.float32)
        self.L.orc_intersect(self.h, _p(rays8), C.c_int64(len(rays8)), _p(ids), _p(tuv))
        return ids, tuv

    def occluded(self, rays8):
        rays8 = np.ascontiguousarray(rays8, np.float32).reshape(-1, 8)
        out = np.zeros(len(rays8), np.uint8)
        self.L.orc_occluded(self.h, _p(rays8), C.c_int64(len(rays8)), _p(out))
        return out

    def bsdf(self, mat, q8):
        q8 = np.ascontiguousarray(q8, np.float32).reshape(-1, 8)
        out = np.zeros((len(q8), 12), np.float32)
        self.L.orc_bsdf(self.h, int(mat), _p(q8), C.c_int64(len(q8)), _p(out))
        return out

    def bdpt_paths(self, rays8, max_splats, seeds=None, injected=None):
        rays8 = np.ascontiguousarray(rays8, np.float32).reshape(-1, 8)
        n = len(rays8)
        li, pix, rgb, nd = np.zeros((n, 3), np.float32), np.zeros((n, max_splats), np.int32), np.zeros((n, max_splats, 3), np.float32), np.zeros(n, np.int32)
        if seeds is not None:
            seeds = np.ascontiguousarray(seeds, np.int32)
            self.L.orc_bdpt_paths(self.h, _p(rays8), _p(seeds), None, C.c_int64(0), C.c_int64(n), int(max_splats), _p(li), _p(pix), _p(rgb), _p(nd))
        else:
            injected = np.ascontiguousarray(injected, np.float32)
            self.L.orc_bdpt_paths(self.h, _p(rays8), None, _p(injected), C.c_int64(injected.shape[1]), C.c_int64(n), int(max_splats), _p(li), _p(pix), _p(rgb), _p(nd))
        return li, pix, rgb, nd

    def render(self, desc=None, threads=8, seed=260450963):
        if desc is not None:
            self.set_config(desc)
        img = np.zeros((self.height, self.width, 3), np.float32)
        sec = self.L.orc_render(self.h, _p(img), int(threads), int(seed))
        return img, float(sec)


def sampler_stream(seed, n):
    out = np.zeros(n, np.float32)
    lib().orc_sampler_stream(int(seed), C.c_int64(n), _p(out))
    return out
