"""Shared test helpers (no dependency on /root/reference at run time)."""
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_obj(path):
    """Minimal Wavefront OBJ reader: triangles with ``v/vt[/vn]`` corners (enough for the reference's
    tests/samples/model.obj).  Returns 0-based numpy arrays."""
    vs, vts, fv, fvt = [], [], [], []
    with open(path) as fh:
        for line in fh:
            t = line.split()
            if not t:
                continue
            if t[0] == "v":
                vs.append([float(x) for x in t[1:4]])
            elif t[0] == "vt":
                vts.append([float(x) for x in t[1:3]])
            elif t[0] == "f":
                corners = [c.split("/") for c in t[1:]]
                assert len(corners) == 3, "triangles only"
                fv.append([int(c[0]) - 1 for c in corners])
                fvt.append([int(c[1]) - 1 if len(c) > 1 and c[1] else 0 for c in corners])
    return dict(vertices=np.asarray(vs, dtype=np.float32), uvs=np.asarray(vts, dtype=np.float32),
                faces=np.asarray(fv, dtype=np.int64), face_uvs_idx=np.asarray(fvt, dtype=np.int64))


_CACHE = {}


def golden(name):
    """Load tests/golden/<name>.npz once."""
    if name not in _CACHE:
        _CACHE[name] = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    return _CACHE[name]
