"""bench.py's contract, as far as it can be checked without a GPU: the reference arm (the unmodified reference on the host cores,
oracle/_ref) prints exactly ONE JSON line on stdout with the keys the driver reads, whatever the libraries print; helper
arithmetic of the render samples."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_reference_arm_prints_one_json_line_with_the_contract_keys():
    from oracle import refprt
    if not refprt.available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "c2", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, f"stdout must carry exactly one line, got {len(lines)}"
    d = json.loads(lines[0])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data",
              "config", "e2e", "cpu_baseline", "rules"):
        assert k in d, k
    assert d["impl"] == "reference" and d["unit"] == "Mrays/s" and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["e2e"] == {"value": d["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert set(d["rules"]) >= {"packet", "single", "headline"} and d["rules"]["headline"] == "packet"
    assert d["value"] > 0 and d["rules"]["single"]["value"] > 0
    assert "workload" in d["config"] and "model" not in d["config"]


def test_stratified_tiles_are_distinct_and_inside_the_frame():
    import bench
    for W, H, count in ((1920, 1080, 128), (3840, 2160, 512), (512, 512, 64)):
        tiles, total = bench.stratified_tiles(W, H, count)
        assert total == ((W + 15) // 16) * ((H + 15) // 16)
        assert len(tiles) == len({tuple(t) for t in tiles}) and 0 < len(tiles) <= count
        for x0, y0, x1, y1 in tiles:
            assert 0 <= x0 < x1 <= W and 0 <= y0 < y1 <= H and x1 - x0 <= 16 and y1 - y0 <= 16


def test_both_arms_describe_the_same_configs():
    import bench
    for wl in ("c1", "c3"):                                       # the procedural scenes of c4 / c5 take seconds to generate
        a, b = bench.render_config(wl), bench.render_config(wl)
        assert a["name"] == b["name"] and a["spp"] == b["spp"] and (a["W"], a["H"]) == (b["W"], b["H"])
    assert bench.c2_config(1, 2, 3, 4, 5) == bench.c2_config(1, 2, 3, 4, 5)
