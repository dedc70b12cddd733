"""bench.py host logic that needs no GPU: both arms print the same `config`, the reference arm runs (tiny workload) with an
explicit thread count, cameras of a step are distinct per rank, and the traffic file of the newest round is picked."""
import json
import os
import subprocess
import sys
import types

import numpy as np

import cases  # noqa: F401  (path setup)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def _args(**kw):
    base = dict(workload="c2", views_per_rank=8, no_balance=False)
    base.update(kw)
    return types.SimpleNamespace(**base)


def test_static_config_is_a_function_of_the_command_line_only():
    a = bench.static_config(_args(), 1)
    b = bench.static_config(_args(), 1)
    assert a == b and a["workload"].startswith("c2:") and a["views_per_rank_per_step"] == 8 and a["global_views_per_step"] == 8
    c = bench.static_config(_args(), 8)
    assert c["global_views_per_step"] == 64 and "dealt to the ranks" in c["cameras"] and "exchange" in c["step"]
    assert "r*8" in bench.static_config(_args(no_balance=True), 8)["cameras"]
    assert not any(k in a for k in ("P_visible", "num_rendered"))           # measured quantities live in workload_stats


def test_cameras_of_a_step_are_distinct_across_ranks():
    cams = [bench.make_workload_camera("c3", i, 64) for i in range(64)]
    centres = np.stack([c.camera_center for c in cams])
    d = np.linalg.norm(centres[:, None] - centres[None], axis=-1) + np.eye(64)
    assert d.min() > 1e-3                                                    # no two views of the step coincide
    # N = 1 keeps round 1's orbit of 8
    assert np.allclose(bench.make_workload_camera("c2", 3, 8).camera_center, bench.make_workload_camera("c2", 3).camera_center)


def test_newest_traffic_prefers_the_highest_round(tmp_path, monkeypatch):
    prof = tmp_path / "profiles"
    prof.mkdir()
    for name, val in (("r01_v5_traffic.json", 1), ("r02_traffic_c2.json", 2), ("r02_traffic_c9.json", 3)):
        (prof / name).write_text(json.dumps({"workload": "c9" if "c9" in name else "c2", "traffic_bytes": {"render_bwd": val}}))
    monkeypatch.setattr(bench, "ROOT", str(tmp_path))
    path, tj = bench.newest_traffic("c2")
    assert os.path.basename(path) == "r02_traffic_c2.json" and tj["traffic_bytes"]["render_bwd"] == 2
    assert bench.newest_traffic("c7") == (None, None)


def test_reference_arm_prints_one_line_with_explicit_threads():
    env = dict(os.environ, OMP_NUM_THREADS="1", RANK="0", WORLD_SIZE="1")           # what torchrun exports
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "c1", "--steps", "1",
                        "--warmup", "0"], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["value"] > 0 and d["cpu_baseline"]["kind"] == "port"
    assert d["cpu_baseline"]["cores"] == bench.host_threads() and "OMP_NUM_THREADS in the environment was '1'" in d["cpu_baseline"]["sample"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["config"] == bench.static_config(_args(workload="c1"), 1)
    # other ranks of a torchrun launch print nothing and exit 0
    r2 = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "c1", "--steps", "1"],
                        env=dict(env, RANK="1", WORLD_SIZE="2"), capture_output=True, text=True, timeout=120)
    assert r2.returncode == 0 and r2.stdout.strip() == ""
