"""bench.py's per-kernel table (host logic only): the two launches of a split kernel fold into one row, the smoother
kernels form one group, and the rows whose B_alg counts loads that pattern tiles skip say so."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def _table():
    prof = {"k_proj4": (6, 0.51), "k_divres4<1>": (6, 0.45), "k_divres4<2>": (6, 0.08),
            "(k_mom3<0, false, true, false>)": (6, 4.4), "k_incr4": (30, 1.0), "k_gs_half_c2<3>": (48, 2.4),
            "k_gs_half_c<3>": (72, 0.5), "k_up4": (30, 1.2), "k_down4<0>": (24, 0.35), "k_down4<1>": (6, 0.38),
            "k_down4<2>": (6, 0.1)}
    blocks = {"k_proj4": 6 * 32768, "k_divres4<1>": 6 * 32768, "k_divres4<2>": 6 * 2660, "k_incr4": 30 * 5000,
              "k_up4": 30 * 5000, "k_down4<0>": 24 * 1200, "k_down4<1>": 6 * 32768, "k_down4<2>": 6 * 2660}
    return bench.kernel_table(prof, blocks, 33554432, 3, 6536.0, {"k_proj4": 1.0e9})


def test_split_launches_are_one_row():
    rows = _table()
    div = [r for r in rows if r["kernel"] == "k_divres4"]
    assert len(div) == 1
    assert abs(div[0]["avg_us"] - (0.45 + 0.08) / 6 * 1e3) < 1e-9
    assert div[0]["launches_per_step"] == 2.0


def test_smoother_group_and_shares():
    rows = _table()
    grp = [r for r in rows if r["kernel"].startswith("smoother group")]
    assert len(grp) == 1 and "k_gs_half_c2" in grp[0]["kernel"] and "k_incr4" in grp[0]["kernel"]
    assert abs(grp[0]["share"] - (1.0 + 2.4 + 0.5) / (0.51 + 0.45 + 0.08 + 4.4 + 1.0 + 2.4 + 0.5 + 1.2 + 0.35 + 0.38 + 0.1)) < 1e-12
    mom = next(r for r in rows if r["kernel"] == "k_mom3")
    assert mom["B_alg"] == 90.0 and abs(mom["frac"] - 90.0 * 33554432 * 6 / 4.4e-3 / 1e9 / 6536.0) < 1e-9
    assert rows == sorted(rows, key=lambda r: -(r["share"] or 0))


def test_pattern_kernels_are_flagged():
    rows = _table()
    for r in rows:
        if r["kernel"].split(" (")[0] in bench.PATTERN_KERNELS and r["frac"] is not None:
            assert r["frac_note"] == bench.PATTERN_NOTE
        else:
            assert "frac_note" not in r
    assert next(r for r in rows if r["kernel"] == "k_proj4")["ncu_dram_bytes_per_launch"] == 1.0e9


def test_unsplit_levels_get_their_own_row():
    down = [r["kernel"] for r in _table() if r["kernel"].startswith("k_down4")]
    assert sorted(down) == ["k_down4", "k_down4 (one launch: levels below the split threshold)"]


def test_base_name():
    assert bench.base_name("(k_mom3<0, false, true>)") == ("k_mom3", ["0", "false", "true"])
    assert bench.base_name("k_up4") == ("k_up4", [])
