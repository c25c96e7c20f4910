"""bench.py's output contract where it can be checked without a GPU: the reference arm (`--impl reference`, the CPU port) prints exactly ONE line
on stdout, valid JSON with the keys the driver reads; a non-zero rank under torchrun prints nothing and exits 0; header constants the Python
binding repeats are the header's."""
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(env_extra, *args):
    env = dict(os.environ, **env_extra)
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, env=env, cwd=ROOT, timeout=600)


def test_reference_arm_prints_one_json_line():
    r = _run({}, "--impl", "reference", "--steps", "1", "--warmup", "0", "--ref-spp", "1")
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.split("\n") if l.strip()]
    assert len(lines) == 1, r.stdout[:500]
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "samples/sec" and d["unit"] == "samples/s" and d["higher_is_better"] is True
    assert d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] == 0 and d["value"] > 0 and d["gpu_launches"] == 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["scene"] == "random" and (d["config"]["width"], d["config"]["height"]) == (600, 400)


def test_reference_arm_other_ranks_exit_quietly():
    r = _run({"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2"}, "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0", "--ref-spp", "1")
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_python_constants_match_header(R):
    h = open(os.path.join(ROOT, "include", "rt_abi.h")).read()
    assert int(re.search(r"#define RT_STATIC_MAX_PIXELS (\d+)u", h).group(1)) == R.STATIC_MAX_PIXELS
    assert int(re.search(r"#define RT_ABI_VERSION (\d+)", h).group(1)) == R.core().rt_abi_version()
    for name, val in (("FAST_MATH", R.FLAG_FAST_MATH), ("STRICT_MATH", R.FLAG_STRICT_MATH), ("COOP", R.FLAG_COOP), ("NO_COOP", R.FLAG_NO_COOP),
                      ("WAVEFRONT", R.FLAG_WAVEFRONT), ("NO_WAVEFRONT", R.FLAG_NO_WAVEFRONT)):
        assert int(re.search(r"#define RT_FLAG_%s (\d+)u" % name, h).group(1)) == val, name
