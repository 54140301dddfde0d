"""Headless CUDA rasteriser (csrc/render.cu): what Interface::render draws (SimulatorCPP/Interface.cpp:55-135) into the image
Camera::get_image returns.  The reference's OpenGL path cannot run on this box, so there is no reference image to compare with
("parity unpinned" for this item): the tests check the drawing rules the reference code states -- colours, painter's order,
camera mapping, flags -- on known geometry."""
import numpy as np
import pytest

from helpers import Golden
from modularevogym_b200 import capi

pytestmark = pytest.mark.gpu

BACKGROUND, GRID, DARK, SOFT, EDGE = (244, 245, 247), (204, 204, 204), (38, 38, 38), (191, 191, 191), (0, 63, 92)
ORANGES = np.array([(255, 245, 235), (254, 230, 206), (253, 208, 162), (253, 174, 107), (253, 141, 60), (241, 105, 19), (217, 72, 1),
                    (166, 54, 3), (127, 39, 4)]) / 255.0
BLUES = np.array([(247, 251, 255), (222, 235, 247), (198, 219, 239), (158, 202, 225), (107, 174, 214), (66, 146, 198), (33, 113, 181),
                  (8, 81, 156), (8, 48, 107)]) / 255.0


def ramp(anchors, row):
    t = row / 99.0 * 8.0
    k = min(int(t), 7)
    c = anchors[k] * (1 - (t - k)) + anchors[k + 1] * (t - k)
    return tuple(int(v) for v in np.floor(c * 255.0 + 0.5))


def has(img, rgb):
    return bool((img.reshape(-1, 3) == np.array(rgb, dtype=np.uint8)).all(axis=1).any())


def pixel_of(world_xy, cam_pos, cam_size, res):
    x = ((world_xy[0] - cam_pos[0]) * 2 / cam_size[0] + 1) * res[0] / 2
    y = (-(world_xy[1] - cam_pos[1]) * 2 / cam_size[1] + 1) * res[1] / 2
    return int(x), int(y)


def test_render_rules_on_walker():
    g = Golden("walker_s0")
    B = capi.Batch(0)
    B.add_world(g.world)
    B.finalize()
    cam_pos, cam_size, res = (0.6, 0.3), (2.4, 1.2), (480, 240)            # 24 x 12 voxels around the robot, 20 px per voxel
    img = B.render(0, cam_pos, cam_size, res)
    assert img.shape == (240, 480, 3) and img.dtype == np.uint8
    for rgb in (BACKGROUND, GRID, DARK, EDGE):
        assert has(img, rgb), rgb
    pos = g["pos0"]
    robot = 0
    boxels = B.object_boxels(0, robot)                                     # cell_type, grid_index, TL,TR,BR,BL, ...
    # at rest the length ratio is 1 up to rounding of the lattice positions: a hair above 1 indexes row int(49.99..) = 49,
    # a hair below (or exactly 1) row 0 + 50 (setColorInterval, Interface.cpp:378-409)
    expect = {1: {DARK}, 2: {SOFT}, 3: {ramp(ORANGES, 49), ramp(ORANGES, 50)}, 4: {ramp(BLUES, 49), ramp(BLUES, 50)}, 5: {DARK}}
    for bx in boxels:
        centre = pos[bx[2:6]].mean(axis=0)
        c, r = pixel_of(centre, cam_pos, cam_size, res)
        assert tuple(int(v) for v in img[r, c]) in expect[int(bx[0])], (bx[:2], tuple(img[r, c]))
    # the ground: voxel (3, 0) of the 100 x 1 strip, centre (0.35, 0.05)
    c, r = pixel_of((0.35, 0.05), cam_pos, cam_size, res)
    assert tuple(img[r, c]) == DARK
    # row 0 is the top of the view: far above the robot there is only background and grid
    top = img[:20].reshape(-1, 3)
    assert set(map(tuple, top.tolist())) <= {BACKGROUND, GRID}
    # a grid line every 20 px: x = 0.6 m (voxel 6) sits in the middle column
    assert tuple(img[5, 240]) == GRID or tuple(img[5, 239]) == GRID
    # flags
    no_grid = B.render(0, cam_pos, cam_size, res, hide_grid=True)
    assert not has(no_grid, GRID) and has(no_grid, DARK)
    no_vox = B.render(0, cam_pos, cam_size, res, hide_boxels=True, hide_edges=True, hide_grid=True, hide_background=True)
    assert (no_vox == 255).all()
    no_edges = B.render(0, cam_pos, cam_size, res, hide_edges=True)
    assert not has(no_edges, EDGE)
    # stepping moves the picture: actuators leave rest length and change colour, the image differs
    acts = np.full((1, B.action_stride), 1.6)
    B.set_actions(acts)
    B.step(3, apply_actions=True, want_flags=False)
    moved = B.render(0, cam_pos, cam_size, res)
    assert (moved != img).any()
    B.close()


def test_viewer_and_camera_of_the_pybind_module():
    """Viewer.render(Camera(True)) + Camera.get_image() as EvoViewer.render('img') uses them (evogym/viewer.py:239-250)"""
    from modularevogym_b200 import simulator_cpp as m
    g = Golden("pusher_s3")
    sim = m.Sim()
    sim.init(g.world.width, g.world.height)
    for o in g.world.load_order():
        add = sim.read_robot_from_array if o.is_robot else sim.read_object_from_array
        assert add(o.grid.astype(np.float64), o.connections.astype(np.float64), o.name, o.x, o.y)
    viewer = m.Viewer(sim)
    viewer.init()
    cam = m.Camera(True)
    cam.set_pos(0.8, 0.3); cam.set_size(2.4, 1.2); cam.set_resolution(240, 120)
    viewer.render(cam, False, False, False, False)
    img = np.array(cam.get_image(), dtype=np.uint8).reshape(cam.get_resolution_height(), cam.get_resolution_width(), 3)
    assert img.shape == (120, 240, 3)
    for rgb in (BACKGROUND, GRID, DARK, SOFT, EDGE):                      # the package is SOFT
        assert has(img, rgb), rgb
    B = capi.Batch(0)
    B.add_world(g.world)
    B.finalize()
    assert np.array_equal(img, B.render(0, (0.8, 0.3), (2.4, 1.2), (240, 120)))
    B.close()
    screen = m.Camera(False)
    screen.set_resolution(64, 48)
    viewer.render(screen)                                                  # no window on this box: nothing to draw, no error
    with pytest.raises(RuntimeError):
        screen.get_image()
