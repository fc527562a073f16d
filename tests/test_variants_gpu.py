"""The run-time A/B switches of the library ("Run-time switches") select slower paths that must produce the
SAME bits as the default: K-MOM by cp.async instead of TMA, pressure kernels without pattern tiles, one-cell-per-thread
pressure kernels, plain launches, the body-free K-MOM variant off, K-DIVRES / K-DOWN as one launch or as two on every level, the half sweep with one or two rows per thread.  The switches are read once per process, so every
variant runs in its own process (tests/variant_worker.py) and the digests of u, p, Δt and pois.n are compared."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SWITCHES = ["WL_NO_TMA", "WL_NO_TILE_FLAGS", "WL_NO_VEC4", "WL_NO_PDL", "WL_NO_NOBODY", "WL_NO_SPLIT", "WL_SPLIT_MIN_TILES", "WL_NO_GS_C2", "WL_GS_C2_MIN_BLOCKS"]


def digest(case, switch=None):
    env = {k: v for k, v in os.environ.items() if k not in SWITCHES}
    if switch:
        env[switch] = "1"            # (WL_SPLIT_MIN_TILES=1, WL_GS_C2_MIN_BLOCKS=1: the large-level forms of K-DIVRES / K-DOWN / the half sweep on every level)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "variant_worker.py"), case], capture_output=True, text=True,
                       timeout=300, env=env, cwd=ROOT)
    lines = [l for l in r.stdout.splitlines() if l.startswith("DIGEST")]
    assert r.returncode == 0 and lines, r.stdout[-2000:] + r.stderr[-2000:]
    return lines[-1].split()[2], lines[-1]


@pytest.mark.parametrize("case", ["sphere", "tgv"])
def test_switches_do_not_change_a_bit(case):
    import __graft_entry__ as ge
    ge.build()
    want, line = digest(case)
    if case == "sphere":                                   # the default run did use all three tile classes
        counts = eval(line.split("]", 1)[1].strip())
        assert all(c > 0 for c in counts), line
    for sw in SWITCHES:
        got, l2 = digest(case, sw)
        assert got == want, (sw, line, l2)
