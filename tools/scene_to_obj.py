"""tools/scene_to_obj.py -- write a flat scene (tests/golden/scene_*.npz) back out as OBJ + MTL so that the
reference binary (oracle/_ref), which only reads OBJ files, can be run where /root/reference does not exist
(the GPU box).  The emitted text is verified to parse back, with the reference-compatible number parser of the
product's loader, into bit-identical arrays; a value that does not round-trip with 9 digits is printed with 17."""
import importlib
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def _fmt(vals, digits):
    return " ".join(("%." + str(digits) + "g") % float(v) for v in vals)


def ppm_lut(maxval=255):
    """the 256 values Tex::pf yields (core.h:439-442): std::pow(float(byte) / maxval, 2.2f) with glibc's powf, the one the reference and
    the product's loader call (numpy's float32 power differs from it by an ulp here and there)"""
    import ctypes
    powf = ctypes.CDLL("libm.so.6").powf
    powf.restype, powf.argtypes = ctypes.c_float, [ctypes.c_float, ctypes.c_float]
    return np.array([powf(float(np.float32(b) / np.float32(maxval)), 2.2) for b in range(256)], np.float32)


def write_ppm(tex, path):
    """inverse of Tex::load (core.h:456-497): gamma-compress to 8 bits (the 256 values pow(b / 255, 2.2f) are strictly increasing, so the
    nearest one is the byte the texel came from) and flip vertically; binary P6"""
    lut = ppm_lut()
    b = np.abs(tex[..., None].astype(np.float64) - lut.astype(np.float64)).argmin(-1).astype(np.uint8)
    h, w, _ = tex.shape
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (w, h))
        f.write(b[::-1].tobytes())


def write_obj(scene, obj_path, digits=9):
    obj_path = Path(obj_path)
    mtl_path = obj_path.with_suffix(".mtl")
    textures = getattr(scene, "textures", [])
    for i, t in enumerate(textures):
        write_ppm(t, obj_path.with_name(f"{obj_path.stem}_tex{i}.ppm"))
    with open(mtl_path, "w") as f:
        for i, m in enumerate(scene.materials):
            f.write(f"newmtl mat{i}\nKd {_fmt(m['kd'], digits)}\nKs {_fmt(m['ks'], digits)}\nKe {_fmt(m['ke'], digits)}\n"
                    f"Tf {_fmt(m['tf'], digits)}\nNs {_fmt([m['ns']], digits)}\nNi {_fmt([m['ni']], digits)}\nillum {int(m['illum'])}\n")
            tex = int(m["tex"]) if "tex" in m.dtype.names else 0
            if tex & 0xffff:
                f.write(f"map_Kd {obj_path.stem}_tex{(tex & 0xffff) - 1}.ppm\n")
            if tex >> 16:
                f.write(f"map_Ks {obj_path.stem}_tex{(tex >> 16) - 1}.ppm\n")
            f.write("\n")
    with open(obj_path, "w") as f:
        f.write(f"mtllib {mtl_path.name}\n")
        for v in scene.positions:
            f.write("v " + _fmt(v, digits) + "\n")
        for n in scene.normals:
            f.write("vn " + _fmt(n, digits) + "\n")
        has_vt = len(getattr(scene, "texcoords", [])) > 0
        if has_vt:
            for t in scene.texcoords:
                f.write("vt " + _fmt(t, digits) + "\n")
        cur_shape, cur_mat = -1, -1
        for t in range(scene.n_tris):
            s, m = int(scene.tri_shape[t]), int(scene.tri_mat[t])
            if s != cur_shape:
                f.write(f"o {scene.shape_names[s] if s < len(scene.shape_names) else 'shape%d' % s}\n")
                cur_shape = s
                cur_mat = -2  # force usemtl after every `o` like Blender does
            if m != cur_mat:
                f.write(f"usemtl mat{m}\n")
                cur_mat = m
            a, b = scene.tri_v[t] + 1, scene.tri_n[t] + 1
            if has_vt and scene.tri_t[t][0] >= 0:
                c = scene.tri_t[t] + 1
                f.write(f"f {a[0]}/{c[0]}/{b[0]} {a[1]}/{c[1]}/{b[1]} {a[2]}/{c[2]}/{b[2]}\n")
            else:
                f.write(f"f {a[0]}//{b[0]} {a[1]}//{b[1]} {a[2]}//{b[2]}\n")
    return obj_path


def export_verified(scene, obj_path):
    trb = importlib.import_module("bidirectional-path-tracing_b200")
    for digits in (9, 17):
        write_obj(scene, obj_path, digits)
        back = trb.HostScene.load_obj(obj_path)
        same = all(np.array_equal(getattr(back, k).view(np.uint8), getattr(scene, k).view(np.uint8))
                   for k in ("positions", "normals", "tri_v", "tri_n", "tri_mat", "tri_shape"))
        mats = all(np.array_equal(back.materials[k], scene.materials[k]) for k in ("kd", "ks", "ke", "tf", "ns", "ni", "illum", "tex"))
        if len(getattr(scene, "textures", [])):
            same = same and np.array_equal(back.texcoords.view(np.uint8), scene.texcoords.view(np.uint8)) and np.array_equal(back.tri_t, scene.tri_t)
            same = same and len(back.textures) == len(scene.textures) and all(np.array_equal(x.view(np.uint32), y.view(np.uint32)) for x, y in zip(back.textures, scene.textures))
        if same and mats:
            return Path(obj_path)
    raise RuntimeError("OBJ export does not round-trip")


if __name__ == "__main__":
    trb = importlib.import_module("bidirectional-path-tracing_b200")
    sc = trb.HostScene.load_npz(sys.argv[1])
    print(export_verified(sc, sys.argv[2]))
