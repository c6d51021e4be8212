"""bench.py's contract that can be checked without a GPU: the reference arm prints exactly one JSON
line on stdout with the agreed keys (everything else goes to stderr)."""
import json
import os
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_json_line(built):
    p = subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"],
                       capture_output=True, text=True, timeout=300, cwd=REPO)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.split("\n") if l.strip()]
    assert len(lines) == 1, p.stdout[:500]
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "edges/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["n_gpus"] == 1 and d["steps"] == 1
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"]


def test_other_ranks_of_the_reference_arm_do_nothing(built):
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    p = subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), "--impl", "reference", "--gpus", "2"], capture_output=True,
                       text=True, timeout=60, cwd=REPO, env=env)
    assert p.returncode == 0 and p.stdout.strip() == ""


import pytest


@pytest.mark.gpu
def test_gpu_arm_prints_one_json_line_with_the_contract_keys(built):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    p = subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), "--steps", "2", "--warmup", "3", "--no-cpu-baseline",
                        "--no-visibility"], capture_output=True, text=True, timeout=600, cwd=REPO)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.split("\n") if l.strip()]
    assert len(lines) == 1, p.stdout[:500]
    d = json.loads(lines[0])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
              "dtype", "data", "config", "e2e", "gpu_launches", "clocks", "roofline", "cpu_baseline"):
        assert k in d, k
    assert d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 3 and d["value"] > 1e9 and d["gpu_launches"] > 0
    assert set(("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step")) <= set(d["e2e"])
    assert d["e2e"]["h2d_bytes_per_step"] == 24_000_000 and 0 < d["e2e"]["value"] < d["value"]
    r = d["roofline"]
    assert r["bound"] in ("hbm", "l2", "issue") and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert set(("sm_mhz", "sm_max_mhz", "reasons")) <= set(d["clocks"])


def test_weak_scaling_regions_line_up_with_the_map_slabs():
    """bench.py --gpus N stacks N S4a regions along z: region g's lattice must sit centred in slab g of the
    tall map (keys [32768 - 256 N + 512 g, +512)), with room for the reach of its rays."""
    import importlib.util
    import numpy as np
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(REPO, "bench.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    from sspp_b200 import scenes
    for regions in (1, 2, 4, 8):
        kmin_z = 32768 - (b.N_VOX // 2) * regions
        for g in range(regions):
            zc = b.region_centre_z(g, regions)
            nodes = scenes.lattice_nodes(4, (b.LATTICE - 1) / 3.0, centre=(0.0, 0.0, zc))  # the lattice's corners and thirds
            keys = np.floor(nodes[:, 2] / b.RES).astype(np.int64) + 32768
            lo, hi = kmin_z + b.N_VOX * g, kmin_z + b.N_VOX * (g + 1)
            reach = int(np.ceil((b.RADIUS + max(b.BOX) / 2) / b.RES)) + 2
            assert keys.min() - reach >= lo and keys.max() + reach < hi
            assert abs((keys.min() - lo) - (hi - 1 - keys.max())) <= 1  # centred in its slab
        cfg = b.workload_config(regions, regions, "peer" if regions > 1 else None)
        assert cfg["nodes"] == regions * 10 ** 6 and cfg["voxels"] == regions * 512 ** 3
