"""bench.py --impl reference: the CPU arm assembles the STATED grid -- Decomp2D blocks with two ghost layers, one oracle per host
thread, whole-grid wall clock.  Checked here on a small grid: the blocks partition the grid, carry the reference's ghost layers,
and a step runs every block's Jacobian + residual."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_cpu_arm_blocks_partition_the_grid_and_run():
    import bench
    from iemic_b200 import grids
    n, m, l = 48, 19, 12
    arm = bench.CpuArm((n, m, l), "global", 8)
    assert (arm.npN, arm.npM) == grids.decomp2d(n, m, 8) == (4, 2)
    owned = 0
    for pm in range(arm.npM):
        for pn in range(arm.npN):
            i0, i1 = arm._split(n, arm.npN, pn)
            j0, j1 = arm._split(m, arm.npM, pm)
            cfg, sub, xs = arm.blocks[pm * arm.npN + pn]
            owned += (i1 - i0) * (j1 - j0)
            # periodic in x with npN > 1: two ghost columns on both sides; in y only towards a neighbour (TRIOS_Domain.C:182-189)
            assert cfg["n"] == (i1 - i0) + 4 and cfg["m"] == (j1 - j0) + 2 and cfg["periodic"] == 0
            assert sub.shape == (l + 2, cfg["m"] + 2, cfg["n"] + 2) and xs.size == 6 * cfg["n"] * cfg["m"] * l
            assert np.all(sub[:, :, 0] == 1) and np.all(sub[:, :, -1] == 1)          # cut sides are walls of the sub-domain
    assert owned == n * m
    assert arm.cells_with_ghosts == sum(b[0]["n"] * b[0]["m"] * l for b in arm.blocks) > n * m * l
    arm.setup()
    sec = arm.step()
    assert sec > 0.0 and all(O is not None for O in arm.oracles)
    # one rank: the block IS the grid, periodic like the global problem, no ghosts
    one = bench.CpuArm((n, m, l), "global", 1)
    cfg1, sub1, xs1 = one.blocks[0]
    assert (cfg1["n"], cfg1["m"], cfg1["periodic"]) == (n, m, 1) and np.array_equal(sub1, one.mask)
    assert "WHOLE grid" in arm.describe() and bench.probe_reference_toolchain()[0].startswith("reference CPU path")
