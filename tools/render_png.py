#!/usr/bin/env python
"""Render one environment of a fixture world with the headless CUDA rasteriser and write a PNG (no image library needed).

    python tools/render_png.py [fixture] [n_steps] [out.png]
"""
import os, struct, sys, zlib
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def write_png(path, img):
    h, w, _ = img.shape
    raw = b"".join(b"\x00" + img[r].tobytes() for r in range(h))
    def chunk(tag, data):
        c = struct.pack(">I", len(data)) + tag + data
        return c + struct.pack(">I", zlib.crc32(tag + data) & 0xffffffff)
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) +
                chunk(b"IDAT", zlib.compress(raw, 9)) + chunk(b"IEND", b""))


def main():
    from helpers import Golden
    from modularevogym_b200 import capi
    name = sys.argv[1] if len(sys.argv) > 1 else "walker_s0"
    n_steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    out = sys.argv[3] if len(sys.argv) > 3 else os.path.join(ROOT, "gpurun_out", f"render_{name}.png")
    g = Golden(name)
    B = capi.Batch(0)
    B.add_world(g.world)
    B.finalize()
    rng = np.random.default_rng(0)
    for _ in range(n_steps):
        B.set_actions(np.clip(rng.uniform(-1, 1, (1, B.action_stride)) + 1, 0.6, 1.6))
        B.step(1, apply_actions=True, want_flags=False)
    robot = B.object_index(0, "robot")
    com = B.object_com(robot)[0]
    img = B.render(0, (com[0] + 0.6, 0.55), (3.2, 1.6), (960, 480))       # EvoViewer's default view is 40 x 20 voxels
    os.makedirs(os.path.dirname(out), exist_ok=True)
    write_png(out, img)
    print(out, img.shape)
    B.close()


if __name__ == "__main__":
    main()
