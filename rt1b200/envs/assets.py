"""Language-Table scene assets: Wavefront OBJ meshes and the URDF files that reference them.

The reference loads 16 coloured block URDFs (``blocks/{red,blue,green,yellow}_{moon,cube,star,pentagon}.urdf``, each naming
one of four OBJ meshes with 105-175 faces), the purple pole, the board (``workspace_real.urdf`` → ``plane.obj``) and the
cylinder effector through PyBullet (``language_table/environments/language_table.py:550-570,648-662``).  PyBullet is not
available here, so this module parses the same files directly — ``v``/``f`` records of the OBJ, ``<mesh filename scale>``,
``<cylinder>``/``<box>`` and ``<color rgba>`` of the URDF — and hands triangle soups to the scene builder.

The asset directory is looked up in this order: ``$RT1B200_ASSET_DIR``, the reference installed for the benchmark arm
(``baseline/_ref/language_table/environments/assets``), the read-only reference checkout.  When none exists the scene
falls back to procedural prisms (``rt1b200.envs.scene.prism``).
"""
from __future__ import annotations

import math
import os
import xml.etree.ElementTree as ET
from functools import lru_cache
from typing import Dict, List, Optional, Tuple

import numpy as np

_REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
_CANDIDATES = (
    os.path.join(_REPO, "baseline", "_ref", "language_table", "environments", "assets"),
    "/root/reference/language_table/environments/assets",
)


def find_asset_dir() -> Optional[str]:
    env = os.environ.get("RT1B200_ASSET_DIR")
    for d in ((env,) if env else ()) + _CANDIDATES:
        if d and os.path.isfile(os.path.join(d, "blocks", "cube.obj")):
            return d
    return None


def load_obj(path: str) -> Tuple[np.ndarray, np.ndarray]:
    """Vertices ``[V,3]`` float32 and triangles ``[F,3]`` int64 of a Wavefront OBJ.  Faces may use ``v``, ``v/vt``,
    ``v//vn`` or ``v/vt/vn`` indices (1-based, negative = relative); polygons are fan-triangulated."""
    verts: List[List[float]] = []
    faces: List[Tuple[int, int, int]] = []
    with open(path, "r") as f:
        for line in f:
            if line.startswith("v "):
                p = line.split()
                verts.append([float(p[1]), float(p[2]), float(p[3])])
            elif line.startswith("f "):
                idx = []
                for tok in line.split()[1:]:
                    i = int(tok.split("/")[0])
                    idx.append(i - 1 if i > 0 else len(verts) + i)
                for k in range(1, len(idx) - 1):
                    faces.append((idx[0], idx[k], idx[k + 1]))
    return np.asarray(verts, np.float32), np.asarray(faces, np.int64).reshape(-1, 3)


def _floats(s: Optional[str], n: int, default: float) -> List[float]:
    if not s:
        return [default] * n
    v = [float(x) for x in s.replace(",", " ").split()]
    return (v + [default] * n)[:n]


def _rpy_matrix(rpy) -> np.ndarray:
    r, p, y = rpy
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
    ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    return rz @ ry @ rx


def _cylinder(radius: float, length: float, sides: int = 16) -> np.ndarray:
    """Triangles of a z-axis cylinder centred at the origin (URDF convention)."""
    a = 2 * math.pi * np.arange(sides) / sides
    ring = np.stack([radius * np.cos(a), radius * np.sin(a)], 1)
    top = np.concatenate([ring, np.full((sides, 1), length / 2)], 1)
    bot = np.concatenate([ring, np.full((sides, 1), -length / 2)], 1)
    ct, cb = np.array([0, 0, length / 2]), np.array([0, 0, -length / 2])
    tris = []
    for i in range(sides):
        j = (i + 1) % sides
        tris += [[ct, top[i], top[j]], [cb, bot[j], bot[i]], [bot[i], bot[j], top[j]], [bot[i], top[j], top[i]]]
    return np.asarray(tris, np.float32)


def _box(size) -> np.ndarray:
    x, y, z = (s / 2 for s in size)
    v = np.array([[-x, -y, -z], [x, -y, -z], [x, y, -z], [-x, y, -z], [-x, -y, z], [x, -y, z], [x, y, z], [-x, y, z]], np.float32)
    quads = [(0, 3, 2, 1), (4, 5, 6, 7), (0, 1, 5, 4), (1, 2, 6, 5), (2, 3, 7, 6), (3, 0, 4, 7)]
    return np.asarray([[v[a], v[b], v[c]] for q in quads for a, b, c in ((q[0], q[1], q[2]), (q[0], q[2], q[3]))], np.float32)


def load_urdf_visuals(path: str) -> List[Dict]:
    """Every ``<visual>`` of a URDF as ``{"tris": [T,3,3] in the robot's base frame, "rgb": (r,g,b) in 0..255}``.
    Fixed joints are followed (their ``<origin>`` chains the child link's frame); other joint types are treated as
    fixed at their zero position, which is all the Language-Table assets need."""
    root = ET.parse(path).getroot()
    base_dir = os.path.dirname(path)
    parent_of: Dict[str, Tuple[str, np.ndarray, np.ndarray]] = {}
    for j in root.findall("joint"):
        o = j.find("origin")
        xyz = np.array(_floats(o.get("xyz") if o is not None else None, 3, 0.0))
        rot = _rpy_matrix(_floats(o.get("rpy") if o is not None else None, 3, 0.0))
        parent_of[j.find("child").get("link")] = (j.find("parent").get("link"), rot, xyz)

    def link_frame(name: str) -> Tuple[np.ndarray, np.ndarray]:
        rot, xyz = np.eye(3), np.zeros(3)
        while name in parent_of:
            name, r, t = parent_of[name]
            rot, xyz = r @ rot, r @ xyz + t
        return rot, xyz

    out = []
    for link in root.findall("link"):
        lrot, lxyz = link_frame(link.get("name"))
        for vis in link.findall("visual"):
            geo = vis.find("geometry")
            if geo is None:
                continue
            o = vis.find("origin")
            vrot = _rpy_matrix(_floats(o.get("rpy") if o is not None else None, 3, 0.0))
            vxyz = np.array(_floats(o.get("xyz") if o is not None else None, 3, 0.0))
            mesh, cyl, box = geo.find("mesh"), geo.find("cylinder"), geo.find("box")
            if mesh is not None:
                fn = os.path.join(base_dir, mesh.get("filename"))
                if not os.path.isfile(fn):
                    continue
                v, f = load_obj(fn)
                tris = (v * np.array(_floats(mesh.get("scale"), 3, 1.0), np.float32))[f]
            elif cyl is not None:
                tris = _cylinder(float(cyl.get("radius")), float(cyl.get("length")))
            elif box is not None:
                tris = _box(_floats(box.get("size"), 3, 1.0))
            else:
                continue
            rot, xyz = lrot @ vrot, lrot @ vxyz + lxyz
            tris = (tris.reshape(-1, 3) @ rot.T + xyz).reshape(-1, 3, 3).astype(np.float32)
            col = vis.find("material/color")
            rgba = _floats(col.get("rgba") if col is not None else None, 4, 1.0)
            out.append({"tris": tris, "rgb": tuple(255.0 * c for c in rgba[:3])})
    return out


# blocks are spawned with Euler (pi/2, 0, yaw) (reference language_table.py:863-866): the meshes are modelled y-up
_BLOCK_UPRIGHT = _rpy_matrix((math.pi / 2, 0.0, 0.0))


@lru_cache(maxsize=None)
def block_mesh(name: str) -> Optional[Tuple[np.ndarray, Tuple[float, float, float]]]:
    """Triangles ``[T,3,3]`` of block ``name`` (e.g. ``red_star``, ``purple_pole``) in its own z-up frame at yaw 0 and
    the URDF colour, or ``None`` when the assets are not installed."""
    d = find_asset_dir()
    if d is None:
        return None
    path = os.path.join(d, "blocks", f"{name}.urdf")
    if not os.path.isfile(path):
        return None
    vis = load_urdf_visuals(path)
    if not vis:
        return None
    tris = np.concatenate([v["tris"] for v in vis], 0)
    tris = (tris.reshape(-1, 3) @ _BLOCK_UPRIGHT.T).reshape(-1, 3, 3).astype(np.float32)
    return tris, vis[0]["rgb"]


@lru_cache(maxsize=None)
def effector_mesh() -> Optional[List[Dict]]:
    """Visuals of the cylinder effector (``suction/cylinder_real.urdf``) in the tool frame (z along the tool axis)."""
    d = find_asset_dir()
    if d is None:
        return None
    path = os.path.join(d, "suction", "cylinder_real.urdf")
    return load_urdf_visuals(path) if os.path.isfile(path) else None


@lru_cache(maxsize=None)
def workspace_mesh() -> Optional[Dict]:
    d = find_asset_dir()
    if d is None:
        return None
    path = os.path.join(d, "workspace_real.urdf")
    vis = load_urdf_visuals(path) if os.path.isfile(path) else []
    return vis[0] if vis else None


def face_normals(tris: np.ndarray) -> np.ndarray:
    n = np.cross(tris[:, 1] - tris[:, 0], tris[:, 2] - tris[:, 0])
    return (n / np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-12)).astype(np.float32)
