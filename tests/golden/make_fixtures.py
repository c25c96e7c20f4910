"""Regenerates the committed fixtures under tests/golden/ from the reference's data files.

Run in the build container only (needs /root/reference and PIL); the GPU box uses the committed outputs.
  ../../ray-tracing_b200/data/earthmap_rgb8.npz
                           : /root/reference/earthmap.jpg decoded to RGB8 (PIL; the reference decodes the same file with
                             JuicyPixels, Scenes.hs:157-165 — decoder differences of +-1 LSB are possible).  It is
                             an INPUT of the product (the image texture of three scenes), so it lives in the package, not here
  published_render_edges.npz : edge maps (uint8) of /root/reference/finalimage.png (600x400) and finalScene400.png (500x500)
  cornellBox1000_box4.npy  : /root/reference/cornellBox1000.png (500x500, the HEAD Main.hs configuration) as 8-bit values
                             box-averaged 4x4 -> float32 [125,125,3]; the loose statistical golden for config 4
"""
import os
import numpy as np
from PIL import Image

here = os.path.dirname(os.path.abspath(__file__))
ref = "/root/reference"
earth = np.asarray(Image.open(os.path.join(ref, "earthmap.jpg")).convert("RGB"), dtype=np.uint8)
np.savez_compressed(os.path.join(here, "..", "..", "ray-tracing_b200", "data", "earthmap_rgb8.npz"), rgb=earth)
png = np.asarray(Image.open(os.path.join(ref, "cornellBox1000.png")).convert("RGB"), dtype=np.float32)
assert png.shape == (500, 500, 3)
box = png.reshape(125, 4, 125, 4, 3).mean(axis=(1, 3)).astype(np.float32)
np.save(os.path.join(here, "cornellBox1000_box4.npy"), box)
# edge maps of the two other published renders (README.md:5,18): |gradient| of the luminance, 3x3 box-blurred, as uint8 — what
# tests/test_oracle.py::test_cameras_and_fixed_geometry_match_published_renders projects the oracle's sphere silhouettes onto
def _edges(path):
    a = np.asarray(Image.open(path).convert("RGB"), dtype=np.float64)
    l = a[..., 0] * 0.299 + a[..., 1] * 0.587 + a[..., 2] * 0.114
    gx = np.zeros_like(l); gy = np.zeros_like(l)
    gx[:, 1:-1] = l[:, 2:] - l[:, :-2]; gy[1:-1] = l[2:] - l[:-2]
    e = np.hypot(gx, gy)
    b = sum(np.roll(np.roll(e, dy, 0), dx, 1) for dy in (-1, 0, 1) for dx in (-1, 0, 1)) / 9.0
    return np.clip(b, 0, 255).astype(np.uint8)
np.savez_compressed(os.path.join(here, "published_render_edges.npz"), finalimage=_edges(os.path.join(ref, "finalimage.png")),
                    finalScene400=_edges(os.path.join(ref, "finalScene400.png")))
print(earth.shape, box.shape)
