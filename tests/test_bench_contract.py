"""bench.py's contract on a machine without a GPU: the reference arm (the CPU oracle port on the host cores)
prints ONE JSON line with the keys the driver reads, ranks other than 0 exit quietly, and the product arm
refuses to run without CUDA instead of falling back to the CPU."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, env=None, timeout=600):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True,
                          timeout=timeout, env=e, cwd=ROOT)


def test_reference_arm_prints_one_json_line_with_the_contract_keys():
    r = _run(["--impl", "reference", "--steps", "1", "--warmup", "1", "--workload", "c1"])
    assert r.returncode == 0, r.stderr
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "it/s" and d["higher_is_better"] is True
    assert d["metric"] == "rasterizer fwd+bwd it/s" and d["value"] > 0 and d["n_gpus"] == 1
    assert d["config"]["workload"].startswith("c1: 640x480, 50000 Gaussians")
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["cpu_baseline"]["value"] == d["value"] == d["e2e"]["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["gpu_launches"] == 0 and d["vs_baseline"] is None and d["dtype"] == "f32" and d["data"] == "synthetic"


def test_reference_arm_honours_steps_and_mirrors_the_mapping_config():
    """N > 1: the reference arm times the mapping workload (one view of the 8-view batch per step) with the
    product arm's config / metric / unit, and runs exactly the steps it was asked for."""
    r = _run(["--impl", "reference", "--gpus", "2", "--steps", "3", "--warmup", "1", "--mapping-workload", "tiny"],
             env={"RANK": "0", "LOCAL_RANK": "0", "WORLD_SIZE": "2"})
    assert r.returncode == 0, r.stderr
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and d["steps"] == 3 and d["warmup"] == 1 and d["n_gpus"] == 2
    assert d["unit"] == "views/s" and d["scaling"] == "strong" and d["metric"].startswith("mapping it/s")
    sys.path[:0] = [ROOT, os.path.join(ROOT, "eags-slam_b200")]
    import bench
    from eags_raster import scene
    assert d["config"] == bench.mapping_config("tiny", scene.workload_camera("tiny"), 300)
    assert "upstream_cuda_rasterizer" in d and d["e2e"]["value"] == d["value"]


def test_reference_arm_other_ranks_exit_without_work():
    r = _run(["--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1"],
             env={"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2"}, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_product_arm_has_no_cpu_fallback():
    r = _run(["--steps", "1", "--warmup", "1", "--no-cpu", "--workload", "c1"], timeout=300)
    assert r.returncode != 0
    assert r.stdout.strip() == ""           # no bench line from a machine that cannot run the kernels
