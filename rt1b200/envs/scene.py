"""Scene description of the Language-Table board: block outlines → prism meshes, the pinhole camera, and projection of
every environment's triangles to screen space.

Replaces the URDF/OBJ loading + PyBullet camera of the reference (``language_table/environments/assets``,
``language_table.py:365-397,550-570,648-662``, ``constants.py:43-58``).  When the reference's asset directory is
installed (``rt1b200.envs.assets.find_asset_dir``) the scene is built from the SAME files: the four block meshes
(cube/moon 112, pentagon 105, star 175 faces) with the colours of the 16 block URDFs, the purple pole, the board plane of
``workspace_real.urdf`` and the cylinder effector of ``suction/cylinder_real.urdf``; faces are flat-shaded (ambient +
Lambert term, the coefficients PyBullet's TinyRenderer defaults to) with the block's current yaw.  Without the assets
the meshes are procedural prisms (square, regular pentagon, five-point star, crescent).  The camera reproduces the
reference's parameters exactly: position (0.75, 0, 0.5), orientation Euler (pi/5, pi, -pi/2) in PyBullet's convention,
look direction R·(0,0,1), up R·(0,-1,0), focal length 0.803·W pixels, 320x180 image.
"""
from __future__ import annotations

import math
from typing import Dict, List, Tuple

import numpy as np
import torch

from language_table.environments import blocks as blocks_module
from language_table.environments import constants as C

from . import assets

TABLE_RGB = (196.0, 180.0, 150.0)
BOARD_RGB = (230.0, 222.0, 205.0)
BACKGROUND_RGB = (90.0, 90.0, 96.0)
EFFECTOR_RGB = (40.0, 40.0, 44.0)
EFFECTOR_SIDES = 12
EFFECTOR_TOP = 0.10
EFFECTOR_CLEARANCE = 0.012                       # lowest point of the effector mesh above the board (asset scene)
LIGHT_DIR = np.array([1.0, 1.0, 1.0]) / math.sqrt(3.0)
LIGHT_AMBIENT, LIGHT_DIFFUSE = 0.6, 0.35          # TinyRenderer's default coefficients (specular left out)


def _regular(n: int, r: float, phase: float = 0.0) -> np.ndarray:
    a = phase + 2 * math.pi * np.arange(n) / n
    return np.stack([r * np.cos(a), r * np.sin(a)], 1)


def outline(shape: str, r: float = C.BLOCK_RADIUS) -> np.ndarray:
    """Counter-clockwise 2-D outline of a block shape (local frame, metres)."""
    if shape == "cube":
        return _regular(4, r * 1.15, math.pi / 4)
    if shape == "pentagon":
        return _regular(5, r * 1.1, math.pi / 2)
    if shape == "star":
        outer, inner = _regular(5, r * 1.25, math.pi / 2), _regular(5, r * 0.55, math.pi / 2 + math.pi / 5)
        return np.stack([outer, inner], 1).reshape(10, 2)
    if shape == "moon":
        k = 7
        a = np.linspace(-0.55 * math.pi, 0.55 * math.pi, k)
        outer = np.stack([r * 1.2 * np.cos(a), r * 1.2 * np.sin(a)], 1)
        inner = np.stack([r * 1.2 * np.cos(a[::-1]) * 0.45 + 0.35 * r, r * 1.05 * np.sin(a[::-1])], 1)
        return np.concatenate([outer, inner], 0) - np.array([0.5 * r, 0.0])
    if shape == "pole":
        return _regular(8, r * 0.6)
    raise ValueError(shape)


def _triangulate(poly: np.ndarray, shape: str) -> List[Tuple[int, int, int]]:
    n = len(poly)
    if shape == "moon":                       # strip between the outer arc (first half) and the inner arc (reversed)
        k = n // 2
        tris = []
        for i in range(k - 1):
            o0, o1, i0, i1 = i, i + 1, n - 1 - i, n - 2 - i
            tris += [(o0, o1, i0), (o1, i1, i0)]
        return tris
    if shape == "star":                       # fan around the centre needs a centre vertex: handled by caller (index n)
        return [(n, i, (i + 1) % n) for i in range(n)]
    return [(0, i, i + 1) for i in range(1, n - 1)]


def prism(shape: str, height: float = C.BLOCK_HEIGHT, r: float = C.BLOCK_RADIUS) -> np.ndarray:
    """Triangles [T, 3, 3] of the extruded outline (top cap + side walls; the bottom is never visible)."""
    poly = outline(shape, r)
    n = len(poly)
    verts = poly
    if shape == "star":
        verts = np.concatenate([poly, np.zeros((1, 2))], 0)
    top = np.concatenate([verts, np.full((len(verts), 1), height)], 1)
    bot = np.concatenate([verts, np.zeros((len(verts), 1))], 1)
    tris = [[top[a], top[b], top[c]] for a, b, c in _triangulate(poly, shape)]
    for i in range(n):
        j = (i + 1) % n
        tris += [[bot[i], bot[j], top[j]], [bot[i], top[j], top[i]]]
    return np.asarray(tris, dtype=np.float32)


def camera() -> Dict[str, np.ndarray]:
    """Camera frame (right, up, forward), position and intrinsics."""
    roll, pitch, yaw = C.CAMERA_ORIENTATION
    cr, sr, cp, sp, cy, sy = math.cos(roll), math.sin(roll), math.cos(pitch), math.sin(pitch), math.cos(yaw), math.sin(yaw)
    rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
    ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    rot = rz @ ry @ rx
    forward = rot @ np.array([0.0, 0.0, 1.0])
    up = rot @ np.array([0.0, -1.0, 0.0])
    right = np.cross(forward, up)
    return {"pos": np.array(C.CAMERA_POSE), "right": right, "up": up, "forward": forward,
            "f": C.CAMERA_INTRINSICS[0], "cx": C.CAMERA_INTRINSICS[2], "cy": C.CAMERA_INTRINSICS[5]}


class SceneBuilder:
    """Builds the per-environment triangle list [E, T, 12] (screen x, y, 1/depth per vertex + RGB) on any device."""

    def __init__(self, block_names: List[str], device="cpu", use_assets=None):
        self.device = torch.device(device)
        self.block_names = list(block_names)
        cam = camera()
        self.cam = {k: torch.as_tensor(v, dtype=torch.float32, device=self.device) for k, v in cam.items()}
        static = []
        board = assets.workspace_mesh() if (assets.find_asset_dir() is not None and use_assets is not False) else None
        # table slab (a little larger than the workspace) and the board surface, both at z = 0 (+epsilon ordering by size)
        for (x0, x1, y0, y1, z, rgb) in ((C.X_MIN - 0.25, C.X_MAX + 0.25, C.Y_MIN - 0.35, C.Y_MAX + 0.35, -0.002, TABLE_RGB),
                                          (C.X_MIN - 0.02, C.X_MAX + 0.02, C.Y_MIN - 0.02, C.Y_MAX + 0.02, 0.0, BOARD_RGB)):
            q = np.array([[x0, y0, z], [x1, y0, z], [x1, y1, z], [x0, y1, z]], np.float32)
            static += [(q[[0, 1, 2]], rgb), (q[[0, 2, 3]], rgb)]
        if board is not None:                       # the board plane of workspace_real.urdf, placed at (0.35, 0, 0)
            bt = board["tris"] + np.array([0.35, 0.0, 0.0], np.float32)
            static = static[:2] + [(t, tuple(LIGHT_AMBIENT * 255.0 * 0 + c for c in board["rgb"])) for t in bt]
        self.static_tris = torch.as_tensor(np.stack([t for t, _ in static]), device=self.device)
        self.static_rgb = torch.as_tensor(np.array([c for _, c in static], np.float32), device=self.device)
        self.use_assets = (assets.find_asset_dir() is not None) if use_assets is None else bool(use_assets)
        if self.use_assets and assets.find_asset_dir() is None:
            raise FileNotFoundError("Language-Table assets requested but no asset directory was found (RT1B200_ASSET_DIR)")
        meshes, colors, owner, normals = [], [], [], []
        for bi, name in enumerate(self.block_names):
            color, shape = name.split("_")
            got = assets.block_mesh(name) if self.use_assets else None
            if got is not None:
                m, rgb = got
                colors.append(np.repeat(np.array(rgb, np.float32)[None], len(m), 0))
                normals.append(assets.face_normals(m))
            else:
                m = prism(shape)
                shade = np.ones((len(m), 1), np.float32)
                shade[-2 * len(outline(shape)):] = 0.72                  # side walls darker than the top cap
                colors.append(np.array(blocks_module.RGB[color], np.float32)[None] * shade)
                normals.append(np.zeros((len(m), 3), np.float32))          # zero normal = colour already final
            meshes.append(m)
            owner += [bi] * len(m)
        self.block_tris = torch.as_tensor(np.concatenate(meshes), device=self.device)          # [Tb, 3, 3] local
        self.block_rgb = torch.as_tensor(np.concatenate(colors), device=self.device)
        self.block_nrm = torch.as_tensor(np.concatenate(normals), device=self.device)          # [Tb, 3] local, or 0
        self.block_owner = torch.as_tensor(owner, device=self.device, dtype=torch.long)
        self.light = torch.as_tensor(LIGHT_DIR, dtype=torch.float32, device=self.device)
        eff_vis = assets.effector_mesh() if self.use_assets else None
        if eff_vis:
            down = np.diag([-1.0, 1.0, -1.0])                             # tool z points at the board (rotvec [0, pi, 0])
            parts = [(v["tris"].reshape(-1, 3) @ down.T).reshape(-1, 3, 3) for v in eff_vis]
            zmin = min(float(p[..., 2].min()) for p in parts)
            eff = np.concatenate(parts).astype(np.float32)
            eff[..., 2] += EFFECTOR_CLEARANCE - zmin
            n = assets.face_normals(eff)
            shade = LIGHT_AMBIENT + LIGHT_DIFFUSE * np.clip(n @ LIGHT_DIR, 0.0, None)           # the effector never rotates
            rgb = np.concatenate([np.repeat(np.array(v["rgb"], np.float32)[None], len(p), 0) for v, p in zip(eff_vis, parts)])
            self.eff_tris = torch.as_tensor(eff, device=self.device)
            self.eff_rgb = torch.as_tensor((rgb * shade[:, None]).astype(np.float32), device=self.device)
        else:
            eff = prism("pole", EFFECTOR_TOP, C.EFFECTOR_RADIUS / 0.6)
            self.eff_tris = torch.as_tensor(eff, device=self.device)
            shade = np.ones((len(eff), 1), np.float32)
            shade[-16:] = 0.75
            self.eff_rgb = torch.as_tensor(np.array(EFFECTOR_RGB, np.float32)[None] * shade, device=self.device)
        self.max_tris = len(self.static_tris) + len(self.block_tris) + len(self.eff_tris)

    def project(self, world: torch.Tensor) -> torch.Tensor:
        """world [..., 3] → screen (x, y, 1/depth)."""
        c = self.cam
        d = world - c["pos"]
        xc, yc, zc = (d * c["right"]).sum(-1), (d * c["up"]).sum(-1), (d * c["forward"]).sum(-1)
        zc = zc.clamp_min(1e-4)
        return torch.stack([c["cx"] + c["f"] * xc / zc, c["cy"] - c["f"] * yc / zc, 1.0 / zc], -1)

    def triangles(self, effector_xy, block_xy, block_yaw, block_mask):
        """effector_xy [E,2], block_xy [E,B,2], block_yaw [E,B], block_mask [E,B] bool → (tris [E,T,12] fp32, ntris [E] int32).
        Triangles of absent blocks are degenerate (zero area) and are skipped by the rasteriser."""
        E = effector_xy.shape[0]
        own = self.block_owner
        cy, sy = torch.cos(block_yaw)[:, own], torch.sin(block_yaw)[:, own]                 # [E,Tb]
        loc = self.block_tris[None].expand(E, -1, -1, -1)                                   # [E,Tb,3,3]
        x = loc[..., 0] * cy[..., None] - loc[..., 1] * sy[..., None] + block_xy[:, own, 0][..., None]
        y = loc[..., 0] * sy[..., None] + loc[..., 1] * cy[..., None] + block_xy[:, own, 1][..., None]
        bw = torch.stack([x, y, loc[..., 2]], -1)
        bw = torch.where(block_mask[:, own][..., None, None], bw, torch.zeros_like(bw))
        ew = self.eff_tris[None].expand(E, -1, -1, -1).clone()
        ew[..., 0] += effector_xy[:, None, None, 0]
        ew[..., 1] += effector_xy[:, None, None, 1]
        world = torch.cat([self.static_tris[None].expand(E, -1, -1, -1), bw, ew], 1)         # [E,T,3,3]
        scr = self.project(world).reshape(E, -1, 9)
        # flat shading with the block's current yaw (faces with a zero normal carry their final colour already)
        nl = self.block_nrm
        nw = torch.stack([nl[:, 0] * cy - nl[:, 1] * sy, nl[:, 0] * sy + nl[:, 1] * cy, nl[:, 2].expand(E, -1)], -1)   # [E,Tb,3]
        lit = LIGHT_AMBIENT + LIGHT_DIFFUSE * (nw * self.light).sum(-1).clamp_min(0.0)
        shade = torch.where((nl.abs().sum(-1) > 0)[None], lit, torch.ones_like(lit))
        brgb = self.block_rgb[None] * shade[..., None]
        rgb = torch.cat([self.static_rgb[None].expand(E, -1, -1), brgb, self.eff_rgb[None].expand(E, -1, -1)], 1)
        tris = torch.cat([scr, rgb], -1).contiguous()
        ntris = torch.full((E,), tris.shape[1], dtype=torch.int32, device=tris.device)
        return tris, ntris
