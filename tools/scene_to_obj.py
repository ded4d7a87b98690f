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


def write_obj(scene, obj_path, digits=9):
    obj_path = Path(obj_path)
    mtl_path = obj_path.with_suffix(".mtl")
    with open(mtl_path, "w") as f:
        for i, m in enumerate(scene.materials):
            f.write(f"newmtl mat{i}\nKd {_fmt(m['kd'], digits)}\nKs {_fmt(m['ks'], digits)}\nKe {_fmt(m['ke'], digits)}\n"
                    f"Tf {_fmt(m['tf'], digits)}\nNs {_fmt([m['ns']], digits)}\nNi {_fmt([m['ni']], digits)}\nillum {int(m['illum'])}\n\n")
    with open(obj_path, "w") as f:
        f.write(f"mtllib {mtl_path.name}\n")
        for v in scene.positions:
            f.write("v " + _fmt(v, digits) + "\n")
        for n in scene.normals:
            f.write("vn " + _fmt(n, digits) + "\n")
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
            f.write(f"f {a[0]}//{b[0]} {a[1]}//{b[1]} {a[2]}//{b[2]}\n")
    return obj_path


def export_verified(scene, obj_path):
    trb = importlib.import_module("bidirectional-path-tracing_b200")
    for digits in (9, 17):
        write_obj(scene, obj_path, digits)
        back = trb.HostScene.load_obj(obj_path)
        same = all(np.array_equal(getattr(back, k).view(np.uint8), getattr(scene, k).view(np.uint8))
                   for k in ("positions", "normals", "tri_v", "tri_n", "tri_mat", "tri_shape"))
        mats = all(np.array_equal(back.materials[k], scene.materials[k]) for k in ("kd", "ks", "ke", "tf", "ns", "ni", "illum"))
        if same and mats:
            return Path(obj_path)
    raise RuntimeError("OBJ export does not round-trip")


if __name__ == "__main__":
    trb = importlib.import_module("bidirectional-path-tracing_b200")
    sc = trb.HostScene.load_npz(sys.argv[1])
    print(export_verified(sc, sys.argv[2]))
