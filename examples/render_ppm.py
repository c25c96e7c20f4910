#!/usr/bin/env python
"""What app/Main.hs does, through the mirrored host API: build a scene, runRender, print the image.

  python examples/render_ppm.py cornell 500 500 1000 out.ppm [--binary]

Main.hs:9-24 fixes scene / size / spp at compile time; here they are arguments.  `run_render` = mkRenderStaticEnv + runRender
(the flattened scene goes through the C ABI to the B200 kernels; there is no CPU path), `write_image` = Main.hs:58-62."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rtb200_loader

pkg = rtb200_loader.load()
from ray_tracing_b200._scene import write_image  # noqa: E402

name, W, H, spp, path = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), sys.argv[5]
rows = pkg.HostScene(name).run_render(W, H, spp, max_depth=50)
write_image(path, rows, binary="--binary" in sys.argv)
print("wrote %s (%dx%d, %d spp)" % (path, W, H, spp), file=sys.stderr)
