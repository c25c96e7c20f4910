"""Progressive accumulation with checkpoint / resume on top of the sample-range form of rt_render.

The reference renders a frame in one go: `renderPos` sums `ns` samples per pixel and divides (Lib.hs:1335-1350), and a run
that is interrupted starts over.  Here sample s of pixel p always draws from streamGen(seed, p, s) (DESIGN.md "RNG"), so a
frame is the sum of any partition of [0, ns) into sample ranges: this class renders the ranges one `step` at a time into a
float64 accumulator, can write the accumulator plus the next sample index to a file, and can pick a frame up from such a
file — in another process or on another GPU — and finish it.  The finished sums equal the single-call sums up to the
order of the double additions (tests: 1e-12 relative).  Plumbing only: every sample is rendered by rt_render.
"""
import json
import os

import numpy as np

_KEYS = ("width", "height", "total_spp", "depth", "seed", "precision", "flags", "tag")


class ProgressiveRender:
    def __init__(self, scene, width, height, total_spp, depth=50, seed=1, precision=1, flags=0, tag=""):  # precision 1 = RT_PREC_DEFAULT (mixed)
        """`scene`: a DeviceScene (anything with render(W, H, spp, depth=, seed=, s0=, precision=, flags=, total=) -> (sums, stats)).
        `tag`: free text naming the scene; stored in checkpoints and compared on resume."""
        if width <= 0 or height <= 0 or total_spp < 0:
            raise ValueError("bad frame size / sample count")
        self.scene = scene
        self.meta = dict(width=int(width), height=int(height), total_spp=int(total_spp), depth=int(depth), seed=int(seed),
                         precision=int(precision), flags=int(flags), tag=str(tag))
        self.sums = np.zeros((height, width, 3), np.float64)
        self.done = 0          # samples per pixel accumulated so far = the next sample index
        self.kernel_ms = 0.0   # device time spent so far (over all processes that worked on this frame)

    @property
    def remaining(self):
        return self.meta["total_spp"] - self.done

    def step(self, n=None):
        """Adds samples [done, done + n) (clipped to the frame's total; n = None: all that remain).  Returns the number rendered."""
        m = self.meta
        n = self.remaining if n is None else min(int(n), self.remaining)
        if n <= 0:
            return 0
        part, st = self.scene.render(m["width"], m["height"], n, depth=m["depth"], seed=m["seed"], s0=self.done,
                                     precision=m["precision"], flags=m["flags"], total=m["total_spp"])
        if st["samples"] != m["width"] * m["height"] * n:
            raise RuntimeError("rt_render traced %d paths, expected %d" % (st["samples"], m["width"] * m["height"] * n))
        self.sums += part
        self.done += n
        self.kernel_ms += st["kernel_ms"]
        return n

    def mean(self):
        """Radiance estimate from the samples so far (renderPos's `divide ns` with ns = done, Lib.hs:1349)."""
        return self.sums / max(self.done, 1)

    def image(self):
        """albedoToColor of the current estimate (device tonemap, Lib.hs:287-292): uint8 [H, W, 3], top row first."""
        from ._scene import tonemap
        return tonemap(self.sums, max(self.done, 1))

    def save(self, path):
        """Checkpoint: accumulator + progress + frame parameters, written atomically (rename over the old file)."""
        tmp = str(path) + ".tmp"
        with open(tmp, "wb") as f:
            np.savez(f, sums=self.sums, state=np.array(json.dumps(dict(self.meta, done=self.done, kernel_ms=self.kernel_ms, format=1))))
        os.replace(tmp, str(path))

    @classmethod
    def resume(cls, scene, path, **expect):
        """Continues the frame stored at `path` on `scene`.  Keyword arguments (width, height, total_spp, depth, seed, precision,
        flags, tag) are checked against the checkpoint: resuming with another seed or size would silently mix two frames."""
        with np.load(str(path)) as z:
            state = json.loads(str(z["state"]))
            sums = np.array(z["sums"], dtype=np.float64)
        if state.get("format") != 1:
            raise ValueError("unknown checkpoint format")
        for k, v in expect.items():
            if k not in _KEYS:
                raise TypeError("unknown frame parameter %r" % k)
            if state[k] != v:
                raise ValueError("checkpoint was made with %s = %r, not %r" % (k, state[k], v))
        self = cls(scene, *(state[k] for k in _KEYS[:3]), **{k: state[k] for k in _KEYS[3:]})
        if sums.shape != self.sums.shape or not 0 <= state["done"] <= state["total_spp"]:
            raise ValueError("corrupt checkpoint")
        self.sums, self.done, self.kernel_ms = sums, int(state["done"]), float(state["kernel_ms"])
        return self
