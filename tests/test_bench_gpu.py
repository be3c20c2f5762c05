"""bench.py's contract on small workloads of every mode, both arms: one JSON line with the keys the driver reads."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*extra):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "3", "--warmup", "3",
                          "--views-per-rank", "2", *extra], capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    return json.loads(lines[0])


@pytest.mark.parametrize("workload", ["smoke_256", "smoke_separate_256", "smoke_forward_256", "smoke_animation_256"])
def test_bench_line_of_every_mode(workload):
    ours = _run("--workload", workload, "--cpu-baseline-seconds", "1")
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "dtype", "data", "config", "e2e", "gpu_launches", "clocks", "roofline", "stages", "cpu_baseline"):
        assert k in ours, k
    assert ours["value"] > 0 and ours["e2e"]["value"] > 0 and ours["gpu_launches"] > 0
    assert ours["roofline"]["bound"] in ("hbm", "issue") and ours["config"]["workload"] == workload
    assert ours["e2e"]["d2h_bytes_per_step"] > 0 and ours["e2e"]["h2d_bytes_per_step"] > 0
    ref = _run("--workload", workload, "--impl", "reference")
    assert ref["impl"] == "reference" and ref["value"] > 0 and ref["reference_device"] == "cuda"
    # same workload description on both arms (the driver compares them)
    for k in ("workload", "n_human", "n_scene", "image", "mode", "global_view_batch"):
        assert ref["config"][k] == ours["config"][k], k
    assert ref["metric"] == ours["metric"] and ref["unit"] == ours["unit"]
