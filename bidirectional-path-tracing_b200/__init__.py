"""bidirectional-path-tracing_b200 -- Python face of the B200-native BDPT hot path.

This package is plumbing only: it loads the in-tree ``libtrb.so`` (host C++ + sm_100a CUDA kernels, built by
``build.py``) and exposes the C ABI of ``include/trb.h`` through ctypes with numpy buffers.  The package
name contains a hyphen, import it with ``importlib.import_module("bidirectional-path-tracing_b200")``.

There is deliberately NO fallback: if the library is missing the import of :func:`lib` raises, and every
compute call fails with ``TrbError`` when no CUDA device is usable.  Nothing under ``oracle/`` is imported
from here (the oracle is test infrastructure).
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

PKG_DIR = Path(__file__).resolve().parent
LIB_PATH = PKG_DIR / "libtrb.so"

TRB_OK, TRB_ERR_INVALID, TRB_ERR_NO_DEVICE, TRB_ERR_CUDA, TRB_ERR_NO_SCENE, TRB_ERR_IO, TRB_ERR_UNSUPPORTED = 0, -1, -2, -3, -4, -5, -6
TRB_FLAG_NO_JITTER = 1
TRB_FLAG_NO_OVERLAP = 2
TRB_FLAG_RUSSIAN_ROULETTE = 4  # the reference's NO_RR 0 behaviour (bdpt.h:18,68,129); default off = shipped behaviour
TRB_RR_DEFAULT_MAX_DEPTH = 32
TRB_FLAG_LIGHT_TRACING = 16  # the reference's estimator A/B switches (bdpt.h:16-17): only the light walk / only the eye walk, no MIS weights
TRB_FLAG_PATH_TRACING = 32
TRB_FLAG_SORT_RAYS = 8      # accepted and ignored (round-1 experiment, removed: profiles/r1f_tuning_variants.md)


class TrbError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"trb error {code}: {msg}")
        self.code = code


# ---- ctypes mirrors of include/trb.h -----------------------------------------------------------------
class Material(C.Structure):
    _fields_ = [("kd", C.c_float * 3), ("ks", C.c_float * 3), ("ke", C.c_float * 3), ("tf", C.c_float * 3),
                ("ns", C.c_float), ("ni", C.c_float), ("illum", C.c_int32), ("tex", C.c_int32)]


class SceneDesc(C.Structure):
    _fields_ = [("positions", C.c_void_p), ("n_positions", C.c_size_t),
                ("normals", C.c_void_p), ("n_normals", C.c_size_t),
                ("tri_v", C.c_void_p), ("tri_n", C.c_void_p), ("tri_mat", C.c_void_p), ("tri_shape", C.c_void_p),
                ("n_tris", C.c_size_t),
                ("materials", C.c_void_p), ("n_materials", C.c_size_t), ("n_shapes", C.c_size_t),
                ("texcoords", C.c_void_p), ("n_texcoords", C.c_size_t), ("tri_t", C.c_void_p),
                ("textures", C.c_void_p), ("n_textures", C.c_size_t)]


class Texture(C.Structure):
    """trb_texture: width*height RGB floats as BitmapTexture3f holds them (gamma-expanded, row 0 = last row of the PPM)"""
    _fields_ = [("rgb", C.c_void_p), ("width", C.c_int32), ("height", C.c_int32)]


class RenderDesc(C.Structure):
    _fields_ = [("eye", C.c_float * 3), ("at", C.c_float * 3), ("up", C.c_float * 3), ("fov_deg", C.c_float),
                ("width", C.c_int32), ("height", C.c_int32), ("spp", C.c_int32), ("rr_depth", C.c_int32),
                ("spp_begin", C.c_int32), ("spp_end", C.c_int32), ("seed", C.c_uint32), ("flags", C.c_uint32),
                ("wave_samples", C.c_int64), ("max_depth", C.c_int32), ("reserved", C.c_int32)]


class PathDesc(C.Structure):
    """trb_path_desc: PathTracerIntegrator settings (main.cpp:94-102 defaults)"""
    _fields_ = [("is_explicit", C.c_int32), ("max_depth", C.c_int32), ("rr_prob", C.c_float), ("emitter_samples", C.c_int32),
                ("bsdf_samples", C.c_int32)]


def path_desc(is_explicit=True, max_depth=-1, rr_prob=0.95, emitter_samples=1, bsdf_samples=0) -> PathDesc:
    return PathDesc(int(bool(is_explicit)), int(max_depth), float(rr_prob), int(emitter_samples), int(bsdf_samples))


TRB_PT_DEFAULT_DEPTH_CAP = 64
TRB_NOALLOC_MAX_RAYS = 1 << 20  # include/trb.h
TRB_UPLOAD_WIDE_COLLAPSE = 0x100


class DirectDesc(C.Structure):
    """trb_direct_desc: DirectIntegrator settings (main.cpp:88-93)"""
    _fields_ = [("strategy", C.c_int32), ("emitter_samples", C.c_int32), ("bsdf_samples", C.c_int32), ("reserved", C.c_int32)]


DI_STRATEGIES = {"mis": 0, "area": 1, "solidAngle": 2, "cosineHemisphere": 3, "bsdf": 4}  # TRB_DI_*; any other name -> -1


def direct_desc(strategy="mis", emitter_samples=1, bsdf_samples=1) -> DirectDesc:
    return DirectDesc(int(DI_STRATEGIES.get(strategy, -1) if isinstance(strategy, str) else strategy), int(emitter_samples), int(bsdf_samples), 0)


class CheckpointInfo(C.Structure):
    """trb_checkpoint_info"""
    _fields_ = [("desc", RenderDesc), ("path", PathDesc), ("integrator", C.c_int32), ("spp_done", C.c_int32), ("scene_hash", C.c_uint64)]


class Stats(C.Structure):
    _fields_ = [("ms_total", C.c_double), ("ms_trace_closest", C.c_double), ("ms_trace_any", C.c_double),
                ("ms_logic", C.c_double), ("samples", C.c_uint64), ("rays_closest", C.c_uint64),
                ("rays_any", C.c_uint64), ("launches", C.c_uint64), ("node_visits", C.c_uint64),
                ("tri_tests", C.c_uint64), ("node_visits_any", C.c_uint64), ("tri_tests_any", C.c_uint64),
                ("retraces", C.c_uint64), ("retraces_any", C.c_uint64), ("rays_deferred", C.c_uint64), ("rays_deferred_any", C.c_uint64),
                ("node_visits_bin", C.c_uint64), ("tri_tests_bin", C.c_uint64), ("node_visits_bin_any", C.c_uint64), ("tri_tests_bin_any", C.c_uint64), ("wave_samples", C.c_uint64)]

    def as_dict(self) -> dict:
        return {k: getattr(self, k) for k, _ in self._fields_}


RAY_DTYPE = np.dtype([("o", "<f4", 3), ("d", "<f4", 3), ("min_t", "<f4"), ("max_t", "<f4")])
HIT_DTYPE = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("hit", "<i4"), ("shape", "<i4"), ("prim", "<i4"),
                      ("tri", "<i4"), ("_pad", "<i4")])
BSDF_QUERY_DTYPE = np.dtype([("wo", "<f4", 3), ("wi", "<f4", 3), ("u", "<f4", 2), ("mat", "<i4"), ("_pad", "<i4")])
BSDF_OUT_DTYPE = np.dtype([("eval", "<f4", 3), ("pdf", "<f4"), ("sample_val", "<f4", 3), ("sample_pdf", "<f4"),
                           ("sample_wi", "<f4", 3), ("type", "<u4")])
MATERIAL_DTYPE = np.dtype([("kd", "<f4", 3), ("ks", "<f4", 3), ("ke", "<f4", 3), ("tf", "<f4", 3), ("ns", "<f4"),
                           ("ni", "<f4"), ("illum", "<i4"), ("tex", "<i4")])
assert RAY_DTYPE.itemsize == 32 and HIT_DTYPE.itemsize == 32 and MATERIAL_DTYPE.itemsize == C.sizeof(Material) == 64
assert BSDF_QUERY_DTYPE.itemsize == 40 and BSDF_OUT_DTYPE.itemsize == 48
# device scene records (csrc/host/tr_host.h), for Context.download
DEV_NODE_DTYPE = np.dtype([("c0min", "<f4", 3), ("c0max", "<f4", 3), ("c1min", "<f4", 3), ("c1max", "<f4", 3), ("ref0", "<i4"), ("ref1", "<i4"), ("pad", "<i4", 2)])
DEV_NODE4_DTYPE = np.dtype([("half", DEV_NODE_DTYPE, 2)])   # two DevNode-shaped halves: children 0,1 and 2,3
assert DEV_NODE4_DTYPE.itemsize == 128
DEV_TRI_DTYPE = np.dtype([("v0", "<f4", 3), ("tri", "<i4"), ("e1", "<f4", 3), ("illk", "<f4"), ("e2", "<f4", 3), ("flags", "<i4")])
DEV_SHADE_DTYPE = np.dtype([("v0", "<f4", 3), ("shape", "<i4"), ("v1", "<f4", 3), ("prim", "<i4"), ("v2", "<f4", 3), ("mat", "<i4"),
                            ("n0", "<f4", 3), ("emitter", "<i4"), ("n1", "<f4", 3), ("tri", "<i4"), ("n2", "<f4", 3), ("pad", "<i4")])
DEV_TRI_UV_DTYPE = np.dtype([("st", "<f4", (3, 2)), ("pad", "<f4", 2)])
assert DEV_NODE_DTYPE.itemsize == 64 and DEV_TRI_DTYPE.itemsize == 48 and DEV_SHADE_DTYPE.itemsize == 96 and DEV_TRI_UV_DTYPE.itemsize == 32

_lib = None


def lib() -> C.CDLL:
    """The loaded libtrb.so. Raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        path = Path(os.environ.get("TRB_LIB", LIB_PATH))
        if not path.exists():
            raise ImportError(f"{path} is missing: run `python bidirectional-path-tracing_b200/build.py` "
                              "(or __graft_entry__.build()); the product has no CPU fallback")
        L = C.CDLL(str(path))
        L.trb_host_last_error.restype = C.c_char_p
        for name in ("trb_last_error", "trb_build_info"):
            if hasattr(L, name):
                getattr(L, name).restype = C.c_char_p
        if hasattr(L, "trb_last_error"):
            L.trb_last_error.argtypes = [C.c_void_p]
        if hasattr(L, "trb_scene_device_bytes"):
            L.trb_scene_device_bytes.restype = C.c_size_t
            L.trb_scene_device_bytes.argtypes = [C.c_void_p]
        _lib = L
    return _lib


def _ptr(a: np.ndarray) -> C.c_void_p:
    return C.c_void_p(a.ctypes.data)


def _chk_host(rc: int) -> int:
    if rc < 0:
        raise TrbError(rc, lib().trb_host_last_error().decode())
    return rc


def make_rays(o, d, min_t=1e-8, max_t=np.finfo(np.float32).max) -> np.ndarray:
    o = np.asarray(o, np.float32).reshape(-1, 3)
    d = np.asarray(d, np.float32).reshape(-1, 3)
    r = np.zeros(len(o), RAY_DTYPE)
    r["o"], r["d"] = o, d
    r["min_t"], r["max_t"] = min_t, max_t
    return r


def render_desc(eye=(0.0, 0.8, 3.8), at=(0.0, 0.8, 0.0), up=(0.0, 1.0, 0.0), fov=30.0, width=800, height=600, spp=1,
                rr_depth=5, seed=260450963, spp_begin=0, spp_end=0, flags=0, wave_samples=0, max_depth=0) -> RenderDesc:
    d = RenderDesc()
    d.eye[:], d.at[:], d.up[:] = eye, at, up
    d.fov_deg = fov
    d.width, d.height, d.spp, d.rr_depth = width, height, spp, rr_depth
    d.spp_begin, d.spp_end, d.seed, d.flags, d.wave_samples = spp_begin, spp_end, seed, flags, wave_samples
    d.max_depth = max_depth
    return d


def max_vertices(desc: RenderDesc) -> int:
    """vertices per subpath: rr_depth-1, or max_depth-1 under TRB_FLAG_RUSSIAN_ROULETTE (trb_abi.cu: max_vertices_of)"""
    if desc.flags & TRB_FLAG_RUSSIAN_ROULETTE:
        md = desc.max_depth if desc.max_depth > 0 else max(TRB_RR_DEFAULT_MAX_DEPTH, desc.rr_depth)
        return max(md - 1, 0)
    return max(desc.rr_depth - 1, 0)


# ---- host scene ------------------------------------------------------------------------------------------
class HostScene:
    """Flat scene arrays (what tinyobj::LoadObj + Scene::load hold). Built from an OBJ file by the library's
    own loader, or from arrays / an .npz fixture."""

    FIELDS = ("positions", "normals", "tri_v", "tri_n", "tri_mat", "tri_shape", "materials")

    def __init__(self, positions, normals, tri_v, tri_n, tri_mat, tri_shape, materials, n_shapes=None, shape_names=None,
                 texcoords=None, tri_t=None, textures=None):
        self.positions = np.ascontiguousarray(positions, np.float32).reshape(-1, 3)
        self.normals = np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
        self.tri_v = np.ascontiguousarray(tri_v, np.int32).reshape(-1, 3)
        self.tri_n = np.ascontiguousarray(tri_n, np.int32).reshape(-1, 3)
        self.tri_mat = np.ascontiguousarray(tri_mat, np.int32).reshape(-1)
        self.tri_shape = np.ascontiguousarray(tri_shape, np.int32).reshape(-1)
        self.materials = np.ascontiguousarray(materials, MATERIAL_DTYPE).reshape(-1)
        self.n_shapes = int(n_shapes if n_shapes is not None else (self.tri_shape.max() + 1 if len(self.tri_shape) else 0))
        self.shape_names = list(shape_names) if shape_names is not None else [f"shape{i}" for i in range(self.n_shapes)]
        # bitmap textures (materials["tex"]): vt array, per-corner vt indices (-1 = none), H x W x 3 float32 images (row 0 = bottom)
        self.texcoords = np.ascontiguousarray(texcoords if texcoords is not None else np.zeros((0, 2)), np.float32).reshape(-1, 2)
        self.tri_t = np.ascontiguousarray(tri_t if tri_t is not None else np.full((len(self.tri_mat), 3), -1), np.int32).reshape(-1, 3)
        self.textures = [np.ascontiguousarray(t, np.float32) for t in (textures or [])]

    @property
    def n_tris(self) -> int:
        return len(self.tri_mat)

    def desc(self) -> SceneDesc:
        d = SceneDesc()
        d.positions, d.n_positions = self.positions.ctypes.data, len(self.positions)
        d.normals, d.n_normals = self.normals.ctypes.data, len(self.normals)
        d.tri_v, d.tri_n = self.tri_v.ctypes.data, self.tri_n.ctypes.data
        d.tri_mat, d.tri_shape = self.tri_mat.ctypes.data, self.tri_shape.ctypes.data
        d.n_tris = self.n_tris
        d.materials, d.n_materials = self.materials.ctypes.data, len(self.materials)
        d.n_shapes = self.n_shapes
        self._tex_views = (Texture * max(len(self.textures), 1))(*[Texture(t.ctypes.data, t.shape[1], t.shape[0]) for t in self.textures])
        d.texcoords, d.n_texcoords, d.tri_t = self.texcoords.ctypes.data, len(self.texcoords), self.tri_t.ctypes.data
        d.textures, d.n_textures = C.cast(self._tex_views, C.c_void_p).value, len(self.textures)
        return d

    @classmethod
    def load_obj(cls, path) -> "HostScene":
        L = lib()
        h = C.c_void_p()
        _chk_host(L.trb_host_load_obj(str(path).encode(), C.byref(h)))
        try:
            d = SceneDesc()
            _chk_host(L.trb_host_scene_desc(h, C.byref(d)))

            def arr(p, n, dt):
                if n == 0:
                    return np.zeros(0, dt)
                return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_char)), shape=(n * np.dtype(dt).itemsize,)).view(dt).copy()

            names = []
            buf = C.create_string_buffer(256)
            for i in range(d.n_shapes):
                _chk_host(L.trb_host_shape_name(h, C.c_size_t(i), buf, C.c_size_t(256)))
                names.append(buf.value.decode(errors="replace"))
            tv = C.cast(d.textures, C.POINTER(Texture))
            textures = [arr(tv[i].rgb, tv[i].width * tv[i].height * 3, np.float32).reshape(tv[i].height, tv[i].width, 3) for i in range(d.n_textures)]
            return cls(arr(d.positions, d.n_positions * 3, np.float32), arr(d.normals, d.n_normals * 3, np.float32),
                       arr(d.tri_v, d.n_tris * 3, np.int32), arr(d.tri_n, d.n_tris * 3, np.int32),
                       arr(d.tri_mat, d.n_tris, np.int32), arr(d.tri_shape, d.n_tris, np.int32),
                       arr(d.materials, d.n_materials, MATERIAL_DTYPE), n_shapes=d.n_shapes, shape_names=names,
                       texcoords=arr(d.texcoords, d.n_texcoords * 2, np.float32),
                       tri_t=arr(d.tri_t, d.n_tris * 3, np.int32) if d.tri_t else None, textures=textures)
        finally:
            L.trb_host_scene_free(h)

    def save_npz(self, path) -> None:
        np.savez_compressed(path, positions=self.positions, normals=self.normals, tri_v=self.tri_v, tri_n=self.tri_n,
                            tri_mat=self.tri_mat, tri_shape=self.tri_shape, materials=self.materials.view(np.uint8),
                            n_shapes=np.int64(self.n_shapes), shape_names=np.array(self.shape_names),
                            **({"texcoords": self.texcoords, "tri_t": self.tri_t, **{f"texture{i}": t for i, t in enumerate(self.textures)}}
                               if self.textures else {}))

    @classmethod
    def load_npz(cls, path) -> "HostScene":
        z = np.load(path, allow_pickle=False)
        return cls(z["positions"], z["normals"], z["tri_v"], z["tri_n"], z["tri_mat"], z["tri_shape"],
                   z["materials"].view(MATERIAL_DTYPE), n_shapes=int(z["n_shapes"]), shape_names=[str(s) for s in z["shape_names"]],
                   texcoords=z["texcoords"] if "texcoords" in z else None, tri_t=z["tri_t"] if "tri_t" in z else None,
                   textures=[z[f"texture{i}"] for i in range(sum(k.startswith("texture") for k in z.files))])

    # host-side derived data (no GPU)
    def bvh_order(self) -> np.ndarray:
        out = np.zeros(self.n_tris, np.int32)
        d = self.desc()
        _chk_host(lib().trb_host_bvh_order(C.byref(d), _ptr(out), C.c_size_t(self.n_tris)))
        return out

    def shape_spheres(self) -> np.ndarray:
        """Scene::getShapeCenter / getShapeRadius per shape (renderer.cpp:294-304, 349-358): (n_shapes, 4) = centre xyz, radius"""
        out = np.zeros((self.n_shapes, 4), np.float32)
        d = self.desc()
        _chk_host(lib().trb_host_shape_spheres(C.byref(d), _ptr(out), C.c_size_t(self.n_shapes)))
        return out

    def emitters(self) -> dict:
        d = self.desc()
        cap = 64
        shape = np.zeros(cap, np.int32)
        area = np.zeros(cap, np.float32)
        rad = np.zeros((cap, 3), np.float32)
        cdf = np.zeros(self.n_tris + 2 * cap, np.float32)
        n = _chk_host(lib().trb_host_emitters(C.byref(d), _ptr(shape), _ptr(area), _ptr(rad), C.c_size_t(cap), _ptr(cdf),
                                              C.c_size_t(len(cdf))))
        return {"n": n, "shape": shape[:n], "area": area[:n], "radiance": rad[:n], "cdf": cdf}


def load_toml(path) -> tuple[RenderDesc, str, str, bool]:
    """loadTOML (src/main.cpp:22-116): returns (render desc, resolved obj path, renderer.type, realtime)."""
    d = RenderDesc()
    obj = C.create_string_buffer(4096)
    typ = C.create_string_buffer(64)
    rt = C.c_int(0)
    _chk_host(lib().trb_host_load_toml(str(path).encode(), C.byref(d), obj, C.c_size_t(4096), typ, C.c_size_t(64), C.byref(rt)))
    return d, obj.value.decode(), typ.value.decode(), bool(rt.value)


def load_toml_path(path) -> PathDesc:
    """the `type = "path"` keys of a scene TOML (src/main.cpp:94-102)"""
    pd = PathDesc()
    _chk_host(lib().trb_host_load_toml_path(str(path).encode(), C.byref(pd)))
    return pd


def load_toml_direct(path) -> DirectDesc:
    """the `type = "direct"` keys of a scene TOML (src/main.cpp:88-93); an unknown samplingStrategy comes back as -1"""
    dd = DirectDesc()
    _chk_host(lib().trb_host_load_toml_direct(str(path).encode(), C.byref(dd)))
    return dd


def save_checkpoint(path, info: CheckpointInfo, rgb: np.ndarray) -> None:
    rgb = np.ascontiguousarray(rgb, np.float32)
    assert rgb.size == info.desc.width * info.desc.height * 3
    _chk_host(lib().trb_host_save_checkpoint(str(path).encode(), C.byref(info), _ptr(rgb)))


def load_checkpoint(path, header_only: bool = False):
    """-> (CheckpointInfo, H x W x 3 float32 | None); raises TrbError for a missing, truncated or corrupted file"""
    info = CheckpointInfo()
    _chk_host(lib().trb_host_load_checkpoint(str(path).encode(), C.byref(info), None, C.c_size_t(0)))
    if header_only:
        return info, None
    rgb = np.zeros((info.desc.height, info.desc.width, 3), np.float32)
    _chk_host(lib().trb_host_load_checkpoint(str(path).encode(), C.byref(info), _ptr(rgb), C.c_size_t(rgb.size)))
    return info, rgb


def checkpoint_matches(a: CheckpointInfo, b: CheckpointInfo) -> bool:
    return bool(lib().trb_host_checkpoint_matches(C.byref(a), C.byref(b)))


def save_png(path, rgb: np.ndarray, exposure: float = 1.0) -> None:
    rgb = np.ascontiguousarray(rgb, np.float32)
    h, w = rgb.shape[:2]
    _chk_host(lib().trb_host_save_png(str(path).encode(), _ptr(rgb), int(w), int(h), C.c_float(exposure)))


def camera_matrices(desc: RenderDesc) -> np.ndarray:
    out = np.zeros((4, 4, 4), np.float32)
    _chk_host(lib().trb_host_camera_matrices(C.byref(desc), _ptr(out)))
    return out


def save_exr(path, rgb: np.ndarray) -> None:
    rgb = np.ascontiguousarray(rgb, np.float32)
    h, w, _ = rgb.shape
    _chk_host(lib().trb_host_save_exr(str(path).encode(), _ptr(rgb), w, h))


# ---- device context --------------------------------------------------------------------------------------
class Context:
    """One GPU + one uploaded scene (trb_context)."""

    def __init__(self, device: int = 0):
        self._L = lib()
        if not hasattr(self._L, "trb_create"):
            raise ImportError("libtrb.so was built without the CUDA half")
        self._h = C.c_void_p()
        rc = self._L.trb_create(int(device), C.byref(self._h))
        if rc < 0:
            raise TrbError(rc, self._L.trb_last_error(None).decode())
        self.scene = None

    def close(self) -> None:
        if self._h:
            self._L.trb_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc: int) -> int:
        if rc < 0:
            raise TrbError(rc, self._L.trb_last_error(self._h).decode())
        return rc

    def upload(self, scene: HostScene, build: str = "auto", wide: str = "sah") -> None:
        """trb_scene_upload_ex: build = "auto" | "host" | "device" - where BVH::build and the triangle gather run (same result);
        wide = "sah" (the SAH upper tree over the reference's leaves, the default) | "collapse" (TRB_UPLOAD_WIDE_COLLAPSE)"""
        d = scene.desc()
        flags = {"auto": 0, "host": 1, "device": 2}[build] | {"sah": 0, "collapse": TRB_UPLOAD_WIDE_COLLAPSE}[wide]
        self._chk(self._L.trb_scene_upload_ex(self._h, C.byref(d), C.c_uint32(flags)))
        self.scene = scene

    def built_on_device(self) -> bool:
        return bool(self._L.trb_scene_built_on_device(self._h))

    def download(self, which: str) -> np.ndarray:
        """test hook (trb_scene_download): "nodes" | "tris" | "shade" | "tri_uv" | "wnodes" of the device scene as structured arrays"""
        idx, dt = {"nodes": (0, DEV_NODE_DTYPE), "tris": (1, DEV_TRI_DTYPE), "shade": (2, DEV_SHADE_DTYPE), "tri_uv": (3, DEV_TRI_UV_DTYPE),
                   "wnodes": (4, DEV_NODE4_DTYPE)}[which]
        n = C.c_size_t(0)
        self._chk(self._L.trb_scene_download(self._h, idx, None, C.c_size_t(0), C.byref(n)))
        out = np.zeros(n.value, dt)
        self._chk(self._L.trb_scene_download(self._h, idx, _ptr(out), C.c_size_t(out.nbytes), C.byref(n)))
        return out

    def refresh_scene(self) -> None:
        """host->device copy of the flattened scene from pinned memory (trb_scene_refresh)"""
        self._chk(self._L.trb_scene_refresh(self._h))

    def scene_bytes(self) -> int:
        return int(self._L.trb_scene_device_bytes(self._h))

    def trace_closest(self, rays: np.ndarray) -> np.ndarray:
        rays = np.ascontiguousarray(rays, RAY_DTYPE)
        out = np.zeros(len(rays), HIT_DTYPE)
        self._chk(self._L.trb_trace_closest(self._h, _ptr(rays), C.c_size_t(len(rays)), _ptr(out)))
        return out

    def trace_any(self, rays: np.ndarray) -> np.ndarray:
        rays = np.ascontiguousarray(rays, RAY_DTYPE)
        out = np.zeros(len(rays), np.uint8)
        self._chk(self._L.trb_trace_any(self._h, _ptr(rays), C.c_size_t(len(rays)), _ptr(out)))
        return out

    def bsdf_batch(self, q: np.ndarray) -> np.ndarray:
        q = np.ascontiguousarray(q, BSDF_QUERY_DTYPE)
        out = np.zeros(len(q), BSDF_OUT_DTYPE)
        self._chk(self._L.trb_bsdf_batch(self._h, _ptr(q), C.c_size_t(len(q)), _ptr(out)))
        return out

    def bdpt_paths(self, desc: RenderDesc, cam_rays: np.ndarray, rng: np.ndarray):
        """BDPTIntegrator::render on given camera rays with injected random numbers (n x stride)."""
        cam_rays = np.ascontiguousarray(cam_rays, RAY_DTYPE)
        rng = np.ascontiguousarray(rng, np.float32)
        n, stride = rng.shape
        assert n == len(cam_rays)
        k = max(max_vertices(desc), 1)
        li = np.zeros((n, 3), np.float32)
        pix = np.full((n, k), -1, np.int32)
        rgb = np.zeros((n, k, 3), np.float32)
        draws = np.zeros(n, np.int32)
        self._chk(self._L.trb_bdpt_paths(self._h, C.byref(desc), _ptr(cam_rays), _ptr(rng), C.c_size_t(n), C.c_size_t(stride),
                                         _ptr(li), _ptr(pix), _ptr(rgb), _ptr(draws)))
        return li, pix, rgb, draws

    def framebuffer_upload(self, rgb: np.ndarray) -> None:
        """restore the context's frame buffer (what accumulate=True adds to), e.g. from a checkpoint"""
        rgb = np.ascontiguousarray(rgb, np.float32)
        self._chk(self._L.trb_framebuffer_upload(self._h, _ptr(rgb), int(rgb.shape[1]), int(rgb.shape[0])))

    def render_path(self, desc: RenderDesc, pd: PathDesc, accumulate: bool = False, download: bool = True):
        """Renderer::render with the PathTracerIntegrator (path.h). Returns (H x W x 3 float32 | None, stats)."""
        st = Stats()
        img = np.zeros((desc.height, desc.width, 3), np.float32) if download else None
        self._chk(self._L.trb_render_path(self._h, C.byref(desc), C.byref(pd), int(accumulate), _ptr(img) if download else None, C.byref(st)))
        return img, st.as_dict()

    def path_paths(self, desc: RenderDesc, pd: PathDesc, cam_rays: np.ndarray, rng: np.ndarray):
        """PathTracerIntegrator::render on given camera rays with injected random numbers (n x stride)."""
        cam_rays = np.ascontiguousarray(cam_rays, RAY_DTYPE)
        rng = np.ascontiguousarray(rng, np.float32)
        n, stride = rng.shape
        assert n == len(cam_rays)
        li = np.zeros((n, 3), np.float32)
        draws = np.zeros(n, np.int32)
        self._chk(self._L.trb_path_paths(self._h, C.byref(desc), C.byref(pd), _ptr(cam_rays), _ptr(rng), C.c_size_t(n), C.c_size_t(stride),
                                         _ptr(li), _ptr(draws)))
        return li, draws

    def render_direct(self, desc: RenderDesc, dd: DirectDesc, accumulate: bool = False, download: bool = True):
        """Renderer::render with the DirectIntegrator (direct.h). Returns (H x W x 3 float32 | None, stats)."""
        st = Stats()
        img = np.zeros((desc.height, desc.width, 3), np.float32) if download else None
        self._chk(self._L.trb_render_direct(self._h, C.byref(desc), C.byref(dd), int(accumulate), _ptr(img) if download else None, C.byref(st)))
        return img, st.as_dict()

    def direct_paths(self, desc: RenderDesc, dd: DirectDesc, cam_rays: np.ndarray, rng: np.ndarray):
        """DirectIntegrator::render on given camera rays with injected random numbers (n x stride)."""
        cam_rays = np.ascontiguousarray(cam_rays, RAY_DTYPE)
        rng = np.ascontiguousarray(rng, np.float32)
        n, stride = rng.shape
        assert n == len(cam_rays)
        li = np.zeros((n, 3), np.float32)
        draws = np.zeros(n, np.int32)
        self._chk(self._L.trb_direct_paths(self._h, C.byref(desc), C.byref(dd), _ptr(cam_rays), _ptr(rng), C.c_size_t(n), C.c_size_t(stride),
                                           _ptr(li), _ptr(draws)))
        return li, draws

    def render_bdpt(self, desc: RenderDesc, accumulate: bool = False, download: bool = True, out=None):
        """Renderer::render (offline branch) with the BDPT integrator. Returns (H x W x 3 float32 | None, stats).
        `out`: the caller's own frame buffer (C-contiguous float32, H x W x 3) - e.g. one in pinned memory, which the library's
        device-to-host copy then reaches at the full speed of the bus; the reference's RenderBuffer is such a caller-owned array."""
        st = Stats()
        if out is not None:
            if out.dtype != np.float32 or not out.flags["C_CONTIGUOUS"] or out.shape != (desc.height, desc.width, 3):
                raise ValueError("out must be a C-contiguous float32 array of shape (height, width, 3)")
            download = True
        img = out if out is not None else (np.zeros((desc.height, desc.width, 3), np.float32) if download else None)
        self._chk(self._L.trb_render_bdpt(self._h, C.byref(desc), int(accumulate), _ptr(img) if download else None, C.byref(st)))
        return img, st.as_dict()

    def render_bdpt_device(self, desc: RenderDesc, d_rgb_ptr: int, accumulate: bool = False) -> dict:
        st = Stats()
        self._chk(self._L.trb_render_bdpt_device(self._h, C.byref(desc), int(accumulate), C.c_void_p(d_rgb_ptr), C.byref(st)))
        return st.as_dict()

    def trace_closest_device(self, d_rays_ptr: int, n: int, d_out_ptr: int) -> float:
        ms = C.c_float(0)
        self._chk(self._L.trb_trace_closest_device(self._h, C.c_void_p(d_rays_ptr), C.c_size_t(n), C.c_void_p(d_out_ptr), C.byref(ms)))
        return ms.value

    def trace_any_device_noalloc(self, d_rays_ptr: int, n: int, d_out_ptr: int) -> None:
        """trb_trace_any_device_noalloc: the occlusion query on device pointers with the library's fixed scratch array (no cudaMalloc)"""
        self._chk(self._L.trb_trace_any_device_noalloc(self._h, C.c_void_p(d_rays_ptr), C.c_size_t(n), C.c_void_p(d_out_ptr)))

    def trace_any_device(self, d_rays_ptr: int, n: int, d_out_ptr: int) -> float:
        ms = C.c_float(0)
        self._chk(self._L.trb_trace_any_device(self._h, C.c_void_p(d_rays_ptr), C.c_size_t(n), C.c_void_p(d_out_ptr), C.byref(ms)))
        return ms.value


def render_bdpt_multi(contexts, desc: RenderDesc):
    """trb_render_bdpt_multi: one frame on the contexts' devices (one context per device, same scene), sample-index sharding and ONE
    ncclReduce inside the library.  Returns (H x W x 3 float32, stats of device 0)."""
    arr = (C.c_void_p * len(contexts))(*[c._h.value for c in contexts])
    st = Stats()
    img = np.zeros((desc.height, desc.width, 3), np.float32)
    contexts[0]._chk(lib().trb_render_bdpt_multi(arr, len(contexts), C.byref(desc), _ptr(img), C.byref(st)))
    return img, st.as_dict()


def render_path_multi(contexts, desc: RenderDesc, pd: PathDesc):
    arr = (C.c_void_p * len(contexts))(*[c._h.value for c in contexts])
    st = Stats()
    img = np.zeros((desc.height, desc.width, 3), np.float32)
    contexts[0]._chk(lib().trb_render_path_multi(arr, len(contexts), C.byref(desc), C.byref(pd), _ptr(img), C.byref(st)))
    return img, st.as_dict()


def render_direct_multi(contexts, desc: RenderDesc, dd: DirectDesc):
    arr = (C.c_void_p * len(contexts))(*[c._h.value for c in contexts])
    st = Stats()
    img = np.zeros((desc.height, desc.width, 3), np.float32)
    contexts[0]._chk(lib().trb_render_direct_multi(arr, len(contexts), C.byref(desc), C.byref(dd), _ptr(img), C.byref(st)))
    return img, st.as_dict()


def build_info() -> str:
    L = lib()
    return L.trb_build_info().decode() if hasattr(L, "trb_build_info") else "host-only"
