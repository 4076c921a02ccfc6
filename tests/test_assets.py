"""Scene assets: the reference's OBJ meshes and URDF files parsed directly (reference
``language_table/environments/language_table.py:550-570,648-662`` loads them through PyBullet)."""
import os

import numpy as np
import pytest
import torch

from rt1b200.envs import assets
from rt1b200.envs.raster import render_reference
from rt1b200.envs.scene import SceneBuilder

needs_assets = pytest.mark.skipif(assets.find_asset_dir() is None, reason="Language-Table assets not installed")


def test_obj_loader_handles_index_forms_and_polygons(tmp_path):
    p = tmp_path / "m.obj"
    p.write_text("# comment\nv 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nv 0 0 1\nvt 0 0\nvn 0 0 1\n"
                 "f 1 2 3\nf 1/1 3/1 4/1\nf 1//1 2//1 5//1\nf 1/1/1 2/1/1 3/1/1 4/1/1\nf -1 -2 -3\n")
    v, f = assets.load_obj(str(p))
    assert v.shape == (5, 3) and f.shape == (6, 3)               # the quad became two triangles
    assert f[3].tolist() == [0, 1, 2] and f[4].tolist() == [0, 2, 3]
    assert f[5].tolist() == [4, 3, 2]                              # negative indices are relative to the end


def test_urdf_visuals_follow_fixed_joints_and_primitives(tmp_path):
    (tmp_path / "tri.obj").write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n")
    (tmp_path / "r.urdf").write_text("""<robot name="r"><link name="a"><visual><origin rpy="0 0 0" xyz="0 0 1"/>
        <geometry><mesh filename="tri.obj" scale="2 2 2"/></geometry><material name="m"><color rgba="1 0.5 0 1"/></material></visual></link>
        <joint name="j" type="fixed"><parent link="a"/><child link="b"/><origin rpy="0 0 0" xyz="0 0 0.5"/></joint>
        <link name="b"><visual><geometry><cylinder length="0.2" radius="0.1"/></geometry></visual></link></robot>""")
    vis = assets.load_urdf_visuals(str(tmp_path / "r.urdf"))
    assert len(vis) == 2
    assert np.allclose(vis[0]["tris"][0], [[0, 0, 1], [2, 0, 1], [0, 2, 1]]) and vis[0]["rgb"] == (255.0, 127.5, 0.0)
    z = vis[1]["tris"][..., 2]
    assert np.isclose(z.min(), 0.4) and np.isclose(z.max(), 0.6)   # cylinder centred on the child link's frame


@needs_assets
def test_reference_block_meshes_have_the_tracked_face_counts():
    d = assets.find_asset_dir()
    faces = {}
    for shape in ("cube", "moon", "pentagon", "star"):
        n_f = sum(1 for line in open(os.path.join(d, "blocks", f"{shape}.obj")) if line.startswith("f "))
        faces[shape] = n_f
    assert faces == {"cube": 112, "moon": 112, "pentagon": 105, "star": 175}
    for name, rgb in (("red_star", (255.0, 102.0, 102.0)), ("blue_cube", (102.0, 102.0, 255.0))):
        tris, col = assets.block_mesh(name)
        assert col == rgb and len(tris) >= faces[name.split("_")[1]]
        pts = tris.reshape(-1, 3)
        assert abs(pts[:, 2].min()) < 1e-6 and np.isclose(pts[:, 2].max(), 0.0381, atol=1e-4)   # upright, 1.5 in tall
        assert np.abs(pts[:, :2]).max() < 0.03


@needs_assets
def test_asset_scene_renders_shaded_blocks():
    names = ["red_moon", "blue_cube", "green_star", "yellow_pentagon"]
    sb = SceneBuilder(names, use_assets=True)
    assert sb.use_assets and sb.max_tris > 700                     # real meshes: hundreds of faces per block
    eff = torch.tensor([[0.3, 0.0]])
    xy = torch.tensor([[[0.25, -0.15], [0.35, 0.1], [0.45, -0.05], [0.5, 0.2]]])
    mask = torch.ones(1, 4, dtype=torch.bool)
    imgs = []
    for yaw in (0.0, 1.0):
        tris, n = sb.triangles(eff, xy, torch.full((1, 4), yaw), mask)
        imgs.append(render_reference(tris.numpy(), n.numpy())[0].astype(np.int32))
    img = imgs[0]
    red = (img[..., 0] > 150) & (img[..., 1] < 120) & (img[..., 2] < 120)
    blue = (img[..., 2] > 150) & (img[..., 0] < 120) & (img[..., 1] < 120)
    assert red.sum() > 20 and blue.sum() > 20
    shades = {tuple(c) for c in img[blue]}
    assert len(shades) >= 2                                         # top and side faces are lit differently
    assert np.abs(imgs[0] - imgs[1]).sum() > 0                      # the yaw turns the mesh and its shading
    fallback = SceneBuilder(names, use_assets=False)
    assert not fallback.use_assets and fallback.max_tris < sb.max_tris
