"""Checkpoint / VTK formats (host side of SURVEY §8f rank 4) on a stand-in simulation that has the mirror's
download / upload / set_dts surface; the device round trip is `test_io_round_trip` in test_parity_gpu.py."""
import os
import re

import numpy as np
import pytest

import waterlily_jl_b200  # noqa: F401  (registers the package under its importable name)
from waterlily_jl_b200 import io as wio


class Flow:
    pass


class StubSim:
    """Host arrays in the mirror's (component, z, y, x) order behind the calls io.py makes."""

    def __init__(self, dims, seed=0):
        rng = np.random.default_rng(seed)
        g = tuple(n + 2 for n in dims[::-1])
        self.L, self.U = 8.0, 2.0
        self.f = dict(p=rng.standard_normal(g).astype(np.float32),
                      u=rng.standard_normal((len(dims),) + g).astype(np.float32))
        self.flow = Flow()
        self.flow.Δt = self._dts = [np.float32(x) for x in (0.25, 0.31, 0.29)]
        self.device_dt = None

    def field_shape(self, name):
        return self.f[name].shape

    def download(self, name):
        return self.f[name].copy()

    def upload(self, name, a):
        assert a.shape == self.f[name].shape
        self.f[name] = np.array(a, np.float32)

    def set_dts(self, dts):
        self._dts[:] = [np.float32(x) for x in dts]
        self.device_dt = float(self._dts[-1])

    def sim_time(self):
        return float(sum(self._dts[:-1])) * self.U / self.L

    def cfl(self):
        return 0.125


@pytest.mark.parametrize("dims", [(6, 5), (6, 5, 4)])
def test_vti_round_trip_and_layout(tmp_path, dims):
    sim = StubSim(dims)
    path = wio.write_vti(str(tmp_path / "a.vti"), {"Velocity": sim.f["u"], "Pressure": sim.f["p"]})
    grid, data = wio.read_vti(path)
    assert grid == sim.f["p"].shape
    assert np.array_equal(data["Pressure"], sim.f["p"]) and np.array_equal(data["Velocity"], sim.f["u"])
    raw = open(path, "rb").read()
    head = raw[:raw.index(b"<AppendedData")].decode()
    n = [d + 2 for d in dims] + [1] * (3 - len(dims))
    # vtk_grid(name, 1:n...) : point extents 0..n-1, origin 1, unit spacing (what the reader ext's size check expects)
    assert 'WholeExtent="%s"' % " ".join(f"0 {k - 1}" for k in n) in head and 'Origin="1 1 1"' in head
    assert re.search(r'Name="Velocity" NumberOfComponents="%d"' % len(dims), head)
    # components first per point, x fastest: the first D floats of the velocity block are u[:, 0, (0,) 0]
    off = int(re.search(r'Name="Velocity"[^>]*offset="(\d+)"', head).group(1))
    blob = raw[raw.index(b"_", raw.index(b"<AppendedData")) + 1:]
    assert int(np.frombuffer(blob[off:off + 8], "<u8")[0]) == sim.f["u"].nbytes
    first = np.frombuffer(blob[off + 8:off + 8 + 4 * 2 * len(dims)], "<f4")
    idx0, idx1 = (0,) * len(dims), (0,) * (len(dims) - 1) + (1,)
    assert np.array_equal(first[:len(dims)], sim.f["u"][(slice(None),) + idx0])
    assert np.array_equal(first[len(dims):], sim.f["u"][(slice(None),) + idx1])


def test_writer_collection_and_restart(tmp_path, monkeypatch):
    """vtkWriter / save! / close and load!(sim, Val(:pvd)) (test/test_ioext.jl:11-26)."""
    monkeypatch.chdir(tmp_path)
    sim = StubSim((6, 5, 4), seed=1)
    w = wio.VTKWriter("run", dir="TEST_DIR")
    w.save(sim)
    sim.f["p"] += 1
    sim.set_dts(list(sim._dts) + [0.27])
    w.save(sim)
    w.close()
    entries = wio._read_pvd("run.pvd")
    assert [f for _, f in entries] == ["TEST_DIR/run_000000.vti", "TEST_DIR/run_000001.vti"]
    assert entries[1][0] == round(sim.sim_time(), 4)
    restart = StubSim((6, 5, 4), seed=2)
    restart.set_dts([0.25])
    w2 = wio.load_pvd(restart, "run.pvd")
    assert np.array_equal(restart.f["p"], sim.f["p"]) and np.array_equal(restart.f["u"], sim.f["u"])
    # Δt[end] = t·L/U, then push!(Δt, CFL)  (ext/WaterLilyReadVTKExt.jl:37-38)
    assert len(restart._dts) == 2 and restart._dts[1] == np.float32(0.125) and restart.device_dt == 0.125
    assert abs(restart.sim_time() - sim.sim_time()) < 1e-3
    assert w2.count == [2] and w2.dir_name == "TEST_DIR" and w2.fname == "run"
    w2.save(restart)
    w2.close()
    assert len(wio._read_pvd("run.pvd")) == 3 and os.path.exists("TEST_DIR/run_000002.vti")
    with pytest.raises(ValueError):
        wio.load_pvd(StubSim((6, 5)), "run.pvd")


def test_npz_checkpoint(tmp_path):
    """save!(fname, sim) / load!(sim; fname) (ext/WaterLilyJLD2Ext.jl:11-49, test/test_ioext.jl:36-47)."""
    sim1, sim2 = StubSim((6, 5, 4), seed=3), StubSim((6, 5, 4), seed=4)
    wio.save("sim1.ckpt", sim1, dir=str(tmp_path))
    wio.load(sim2, fname="sim1.ckpt", dir=str(tmp_path))
    assert np.array_equal(sim1.f["p"], sim2.f["p"]) and np.array_equal(sim1.f["u"], sim2.f["u"])
    assert sim1._dts == sim2._dts and sim2.device_dt == float(sim1._dts[-1])
    with pytest.raises(ValueError):
        wio.load(StubSim((6, 5, 6)), fname="sim1.ckpt", dir=str(tmp_path))


def test_slab_pieces_reassemble(tmp_path, monkeypatch):
    """Slab-parallel output: each rank writes its slab as one piece of the whole extent and rank 0 the .pvti index;
    read back, the pieces give the global arrays (adjacent pieces share one point plane)."""
    monkeypatch.chdir(tmp_path)
    nranks, dims = 2, (6, 5, 8)
    glob = StubSim(dims, seed=7)
    w = [wio.VTKWriter("slab", dir="OUT") for _ in range(nranks)]
    for r in range(nranks):
        s = StubSim(dims, seed=7)
        s.rank, s.nranks, s.dims = r, nranks, dims
        k0, n = waterlily_jl_b200.slab_range(r, nranks, dims[-1])
        s.f = {k: np.ascontiguousarray(v[..., k0:k0 + n + 2, :, :]) for k, v in glob.f.items()}
        path = w[r].save(s)
        w[r].close(rank=r)
    assert path == "OUT/slab_000000.pvti" and sorted(os.listdir("OUT")) == ["slab_000000.pvti", "slab_000000_r000.vti", "slab_000000_r001.vti"]
    assert wio.slab_piece(0, 2, 8) == (0, 6, 0) and wio.slab_piece(1, 2, 8) == (5, 5, 1)
    grid, data = wio.read_pvti(path)
    assert grid == glob.f["p"].shape
    assert np.array_equal(data["Pressure"], glob.f["p"]) and np.array_equal(data["Velocity"], glob.f["u"])
    assert [f for _, f in wio._read_pvd("slab.pvd")] == ["OUT/slab_000000.pvti"]
