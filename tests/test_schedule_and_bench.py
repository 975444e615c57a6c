"""Host-side pieces of the step schedules and the bench contract that need no GPU."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _paint(shape, boxes):
    a = np.zeros(shape, np.int32)  # [k][j][i]
    for lo, hi in boxes:
        a[lo[2]:hi[2], lo[1]:hi[1], lo[0]:hi[0]] += 1
    return a


@pytest.mark.parametrize("g", [(6, 7, 8), (66, 66, 66), (18, 18, 18), (5, 5, 5)])
def test_shell_slabs_tile_the_grid_minus_the_inner_box(built, g):
    """the boxes of the library's sweep schedule (bricks_step_boxes; weak/main.cu:246-287 has no such
    split): shell = every cell outside the inner box exactly once, none inside; owned skin = owned
    box minus inner box"""
    inner = built.step_boxes(g, built.PLAN_INNER)
    assert inner == [((2, 2, 2), tuple(x - 2 for x in g))]
    (inner_lo, inner_hi), = inner
    assert built.step_boxes(g, built.PLAN_FULL) == [((0, 0, 0), g)]
    assert built.step_boxes(g, built.PLAN_OWNED) == [((1, 1, 1), tuple(x - 1 for x in g))]
    a = _paint(g[::-1], built.step_boxes(g, built.PLAN_SHELL))
    want = np.ones(g[::-1], np.int32)
    want[inner_lo[2]:inner_hi[2], inner_lo[1]:inner_hi[1], inner_lo[0]:inner_hi[0]] = 0
    assert np.array_equal(a, want)
    # owned skin = owned box minus inner box (the bricks the next exchange sends)
    b = _paint(g[::-1], built.step_boxes(g, built.PLAN_OSKIN))
    want2 = np.zeros(g[::-1], np.int32)
    want2[1:g[2] - 1, 1:g[1] - 1, 1:g[0] - 1] = 1
    want2[inner_lo[2]:inner_hi[2], inner_lo[1]:inner_hi[1], inner_lo[0]:inner_hi[0]] = 0
    assert np.array_equal(b, want2)


def test_subdomains_without_an_inner_box_have_no_split_plans(built):
    for which in (built.PLAN_INNER, built.PLAN_SHELL, built.PLAN_OSKIN):
        assert built.step_boxes((4, 4, 4), which) == []
    assert built.step_boxes((4, 4, 4), built.PLAN_OWNED) == [((1, 1, 1), (3, 3, 3))]


def test_owned_skin_covers_every_brick_an_exchange_sends(built):
    """the early-exchange schedule may start the exchange after the owned-skin slabs only if
    every skin region of BrickDecomp lies inside them"""
    dec = built.Decomp((48, 32, 40), 8)
    g = tuple(int(x) for x in dec.gdims)
    slabs = built.step_boxes(g, built.PLAN_OSKIN)
    covered = _paint(g[::-1], slabs).astype(bool)
    ids = set(int(x) for x in dec.grid[covered])
    for r in dec.skin:
        for b in range(int(r["pos"]), int(r["pos"]) + int(r["len"])):
            assert b in ids
    dec.close()


def test_interior_points_matches_the_reference_count(built):
    from bricklib_b200 import weak
    assert weak.interior_points((512, 512, 512), 8) == 8 * 512 ** 3


def test_reference_arm_prints_the_contract_line(built, oracle):
    """bench.py --impl reference: the reference's CPU path on this arm's metric/config"""
    from oracle import refload
    if not refload.available():
        pytest.skip("oracle/_ref is not built")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1",
                        "--workload", "weak7"], capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stderr[-2000:]
    line = json.loads(p.stdout.strip().splitlines()[-1])
    for key in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
                "scaling", "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert key in line, key
    assert line["impl"] == "reference" and line["unit"] == "GStencil/s" and line["dtype"] == "f64"
    assert line["value"] > 0 and line["e2e"]["h2d_bytes_per_step"] == 0 and line["cpu_baseline"]["kind"] == "reference"
    assert "7-point" in line["config"]["workload"]


def test_bench_refuses_to_run_the_gpu_arm_without_a_gpu(built):
    import ctypes
    n = ctypes.c_int()
    if built.lib().bricks_device_count(ctypes.byref(n)) == 0 and n.value > 0:
        pytest.skip("a GPU is present")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1"], capture_output=True, text=True,
                       timeout=600)
    assert p.returncode != 0 and "no CPU fallback" in (p.stderr + p.stdout)
