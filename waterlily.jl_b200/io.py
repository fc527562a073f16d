"""Checkpoint and output formats for a device-resident simulation.

Host-side only: every function moves whole fields through `Simulation.download` / `upload`, exactly as the
reference's extensions work on `Array(flow.p)` / `copyto!(flow.p, ...)`.

* `save` / `load`                 : `save!(fname, sim)` / `load!(sim; fname)`   (ext/WaterLilyJLD2Ext.jl:11-49).
  The container is `.npz` instead of JLD2 (no HDF5 writer in this image); the payload is the same three
  entries `p`, `u` and the Δt vector (stored under the key `dt`), arrays in the mirror's own (component, z, y, x)
  order.  On a slab-decomposed run every rank stores / restores ITS slab in `fname.rNNNofMMM`.
* These are the mirror's OWN formats: `read_vti` parses what `write_vti` writes (uncompressed raw-appended ImageData) and
  rejects anything else with a clear error; files written by the reference's JLD2 / WriteVTK extensions are read by
  those extensions, which work unchanged on `B200Array` (Julia side).
* `save_meanflow` / `load_meanflow`: the `MeanFlow` pair of the same extension (:24-31, :57-69).
* `VTKWriter` + `load_pvd`        : `vtkWriter` / `save!(w, sim)` / `close(w)` (ext/WaterLilyWriteVTKExt.jl:27-71) and
  `load!(sim, Val(:pvd))` (ext/WaterLilyReadVTKExt.jl:22-43): one ImageData `.vti` per save on the `1:n` point grid
  WriteVTK's `vtk_grid(name, 1:n...)` produces, vectors stored components-first per point, and a `.pvd` collection.

Julia users keep using the reference's own extensions: they only need `Array(::B200Array)` and
`copyto!(::B200Array, ::Array)`, which the shim provides.
"""
import os
import re

import numpy as np

__all__ = ["save", "load", "save_meanflow", "load_meanflow", "VTKWriter", "load_pvd", "write_vti", "read_vti",
           "write_pvti", "read_pvti", "slab_piece", "default_attrib"]


# ------------------------------------------------------------------------------------------------ JLD2 pair
def _slab_name(fname, sim):
    n = getattr(sim, "nranks", 1)
    return fname if n == 1 else "%s.r%03iof%03i" % (fname, sim.rank, n)


def save(fname, sim, dir="./"):
    """`save!(fname, sim; dir)`: pressure, velocity and the Δt vector (every rank its own slab file on several GPUs)."""
    path = os.path.join(dir, _slab_name(fname, sim))
    with open(path, "wb") as fh:                       # a file handle keeps numpy from appending ".npz"
        np.savez(fh, p=sim.download("p"), u=sim.download("u"), dt=np.asarray(sim.flow.Δt, np.float32))
    return path


def load(sim, fname="WaterLily.npz", dir="./"):
    """`load!(sim; fname, dir)`: p and u are overwritten, Δt is replaced (empty! + push!)."""
    with np.load(os.path.join(dir, _slab_name(fname, sim))) as obj:
        p, u, dt = obj["p"], obj["u"], obj["dt"]
    if tuple(p.shape) != tuple(sim.field_shape("p")):
        raise ValueError("Simulation size does not match the size of the stored simulation.")
    sim.upload("p", p)
    sim.upload("u", u)
    sim.set_dts(dt)


def save_meanflow(fname, meanflow, dir="./"):
    """`save!(fname, meanflow; dir)`: P, U, UU (absent without uu_stats) and the time vector."""
    path = os.path.join(dir, fname)
    data = dict(P=meanflow.P, U=meanflow.U, t=np.asarray(meanflow.t, np.float32))
    if meanflow.uu_stats:
        data["UU"] = meanflow.UU
    with open(path, "wb") as fh:
        np.savez(fh, **data)
    return path


def load_meanflow(meanflow, fname="WaterLilyMean.npz", dir="./"):
    """`load!(meanflow; fname, dir)`."""
    with np.load(os.path.join(dir, fname)) as obj:
        data = {k: obj[k] for k in obj.files}
    sim = meanflow.sim
    if tuple(data["P"].shape) != tuple(sim.field_shape("mean_P")):
        raise ValueError("MeanFlow size does not match the size of the stored MeanFlow.")
    sim.upload("mean_P", data["P"])
    sim.upload("mean_U", data["U"])
    if meanflow.uu_stats and "UU" in data:
        sim.upload("mean_UU", data["UU"])
    meanflow.t[:] = [np.float32(x) for x in data["t"]]


# ------------------------------------------------------------------------------------------------ .vti files
def write_vti(path, point_data, grid=None, whole_nz=None, k0=0):
    """One ImageData file on the point grid 1:nx × 1:ny (× 1:nz).  `point_data` maps a name to an array shaped
    `grid` = (nz, ny, nx) / (ny, nx) for scalars or (ncomp,) + grid for vectors; vectors are written components-first
    per point (`components_first`, ext/WaterLilyWriteVTKExt.jl:71).  Raw appended binary, UInt64 block headers, no
    compression.  Without `grid` the lowest-rank array defines it.  `whole_nz`/`k0` place the file as one piece
    (planes k0 .. k0+nz-1) of a taller whole extent: the slab of one rank (see `write_pvti`)."""
    point_data = {k: np.ascontiguousarray(a, np.float32) for k, a in point_data.items()}
    if grid is None:
        grid = min((a.shape for a in point_data.values()), key=len)
    grid = tuple(int(g) for g in grid)
    arrays = []
    for name, a in point_data.items():
        if a.shape == grid:
            ncomp, flat = 1, a.ravel()
        elif a.shape[1:] == grid:
            ncomp, flat = a.shape[0], np.moveaxis(a, 0, -1).ravel()
        else:
            raise ValueError(f"{name}: shape {a.shape} does not fit the grid {grid}")
        arrays.append((name, ncomp, flat))
    n = list(grid[::-1]) + [1] * (3 - len(grid))              # nx, ny, nz
    extent = " ".join(f"0 {k - 1}" for k in n)
    piece = extent
    if whole_nz is not None:
        piece = f"0 {n[0] - 1} 0 {n[1] - 1} {k0} {k0 + n[2] - 1}"
        extent = f"0 {n[0] - 1} 0 {n[1] - 1} 0 {whole_nz - 1}"
    head = ['<?xml version="1.0"?>',
            '<VTKFile type="ImageData" version="1.0" byte_order="LittleEndian" header_type="UInt64">',
            f'  <ImageData WholeExtent="{extent}" Origin="1 1 1" Spacing="1 1 1">',
            f'    <Piece Extent="{piece}">', '      <PointData>']
    offset = 0
    for name, ncomp, flat in arrays:
        head.append(f'        <DataArray type="Float32" Name="{name}" NumberOfComponents="{ncomp}" '
                    f'format="appended" offset="{offset}"/>')
        offset += 8 + flat.nbytes
    head += ['      </PointData>', '    </Piece>', '  </ImageData>', '  <AppendedData encoding="raw">', '_']
    with open(path, "wb") as fh:
        fh.write("\n".join(head).encode())
        for _, _, flat in arrays:
            fh.write(np.uint64(flat.nbytes).tobytes())
            fh.write(flat.astype("<f4", copy=False).tobytes())
        fh.write(b"\n  </AppendedData>\n</VTKFile>\n")
    return path


def read_vti(path):
    """Inverse of `write_vti`: (grid shape as (nz, ny, nx) or (ny, nx), {name: array}) with vectors components-first
    in the leading axis again (`components_last` of the reader ext, in this mirror's axis order)."""
    raw = open(path, "rb").read()
    if b'<AppendedData encoding="raw">' not in raw or b"compressor=" in raw[:4096]:
        raise ValueError(f"{path}: only the uncompressed raw-appended ImageData layout written by write_vti is supported "
                         "(files from WriteVTK.jl are compressed / base64: read them with the reference's ReadVTK extension)")
    mark = raw.index(b'<AppendedData encoding="raw">')
    head = raw[:mark].decode()
    blob = raw[raw.index(b"_", mark) + 1:]
    ext = [int(x) for x in re.search(r'<Piece Extent="([^"]+)"', head).group(1).split()]
    n = [ext[2 * k + 1] - ext[2 * k] + 1 for k in range(3)]   # nx, ny, nz
    grid = tuple(n[::-1]) if n[2] > 1 else (n[1], n[0])
    out = {}
    for m in re.finditer(r'<DataArray type="Float32" Name="([^"]+)" NumberOfComponents="(\d+)" format="appended" offset="(\d+)"/>', head):
        name, ncomp, off = m.group(1), int(m.group(2)), int(m.group(3))
        nbytes = int(np.frombuffer(blob[off:off + 8], "<u8")[0])
        a = np.frombuffer(blob[off + 8:off + 8 + nbytes], "<f4")
        out[name] = a.reshape(grid).copy() if ncomp == 1 else np.moveaxis(a.reshape(grid + (ncomp,)), -1, 0).copy()
    return grid, out


def slab_piece(rank, nranks, nz_global):
    """Point planes of the GLOBAL grid (0 .. nz_global+1, ghosts included) that rank `rank` writes, as
    (first plane, number of planes, planes to drop at the start of its local array).  Every rank holds its interior
    planes plus one ghost plane per side; neighbouring pieces share one point plane (the upper ghost plane of the lower
    rank), which parallel ImageData needs to leave no gap between pieces."""
    n = nz_global // nranks
    k_off = rank * n
    drop = 0 if rank == 0 else 1
    return k_off + drop, n + 2 - drop, drop


def write_pvti(path, pieces, names, nx, ny, whole_nz):
    """Parallel ImageData index: `pieces` = [(file, k0, nz)], `names` = {array name: ncomp}."""
    ext = f"0 {nx - 1} 0 {ny - 1} 0 {whole_nz - 1}"
    lines = ['<?xml version="1.0"?>', '<VTKFile type="PImageData" version="1.0" byte_order="LittleEndian" header_type="UInt64">',
             f'  <PImageData WholeExtent="{ext}" GhostLevel="0" Origin="1 1 1" Spacing="1 1 1">', '    <PPointData>']
    lines += [f'      <PDataArray type="Float32" Name="{k}" NumberOfComponents="{c}"/>' for k, c in names.items()]
    lines += ['    </PPointData>']
    lines += [f'    <Piece Extent="0 {nx - 1} 0 {ny - 1} {k0} {k0 + nz - 1}" Source="{os.path.basename(f)}"/>' for f, k0, nz in pieces]
    lines += ['  </PImageData>', '</VTKFile>', '']
    with open(path, "w") as fh:
        fh.write("\n".join(lines))
    return path


def read_pvti(path):
    """Assemble the global arrays from a `.pvti` and its pieces: (grid, {name: array})."""
    text = open(path).read()
    ext = [int(x) for x in re.search(r'WholeExtent="([^"]+)"', text).group(1).split()]
    grid = (ext[5] + 1, ext[3] + 1, ext[1] + 1)
    out = {}
    for m in re.finditer(r'<Piece Extent="([^"]+)" Source="([^"]+)"', text):
        e = [int(x) for x in m.group(1).split()]
        _, data = read_vti(os.path.join(os.path.dirname(path), m.group(2)))
        for name, a in data.items():
            full = out.setdefault(name, np.zeros(a.shape[:-3] + grid, np.float32))
            full[..., e[4]:e[5] + 1, :, :] = a
    return grid, out


# ------------------------------------------------------------------------------------------------ writer + pvd
def default_attrib():
    """`default_attrib()` (ext/WaterLilyWriteVTKExt.jl:44-46): Velocity and Pressure."""
    return {"Velocity": lambda sim: sim.download("u"), "Pressure": lambda sim: sim.download("p")}


def _read_pvd(fname):
    text = open(fname).read()
    return [(float(t), f) for t, f in re.findall(r'<DataSet timestep="([^"]+)"[^>]*file="([^"]+)"', text)]


class VTKWriter:
    """`vtkWriter(fname; attrib, dir)` (ext/WaterLilyWriteVTKExt.jl:27-35).  `save(sim)` writes `dir/fname_%06i.vti` and
    records `round(sim_time(sim), digits=4)`; `close()` writes `fname.pvd`."""

    def __init__(self, fname="WaterLily", attrib=None, dir="vtk_data", entries=None, count=0):
        os.makedirs(dir, exist_ok=True)
        self.fname, self.dir_name = fname, dir
        self.output_attrib = attrib if attrib is not None else default_attrib()
        self.entries = list(entries or [])
        self.count = [count]

    def save(self, sim):
        k = self.count[0]
        data = {name: np.asarray(fn(sim), np.float32) for name, fn in self.output_attrib.items()}
        grid = tuple(sim.field_shape("p"))
        nranks = getattr(sim, "nranks", 1)
        if nranks == 1:
            path = os.path.join(self.dir_name, "%s_%06i.vti" % (self.fname, k))
            write_vti(path, data, grid=grid)
        else:
            # slab decomposition: every rank writes its own piece, nothing is gathered;
            # rank 0 adds the .pvti index and owns the collection file
            nz_g = sim.dims[-1]
            k0, nz, drop = slab_piece(sim.rank, nranks, nz_g)
            piece = os.path.join(self.dir_name, "%s_%06i_r%03i.vti" % (self.fname, k, sim.rank))
            write_vti(piece, {n_: a[..., drop:, :, :] for n_, a in data.items()}, grid=(nz,) + grid[1:], whole_nz=nz_g + 2, k0=k0)
            path = os.path.join(self.dir_name, "%s_%06i.pvti" % (self.fname, k))
            if sim.rank == 0:
                pieces = [(os.path.join(self.dir_name, "%s_%06i_r%03i.vti" % (self.fname, k, r)),) + slab_piece(r, nranks, nz_g)[:2]
                          for r in range(nranks)]
                write_pvti(path, pieces, {n_: (1 if a.shape == grid else a.shape[0]) for n_, a in data.items()},
                           grid[-1], grid[-2], nz_g + 2)
        self.count[0] = k + 1
        self.entries.append((round(float(sim.sim_time()), 4), path))
        return path

    def close(self, rank=0):
        if rank != 0:                                      # slab runs: rank 0 writes the collection
            return
        lines = ['<?xml version="1.0"?>', '<VTKFile type="Collection" version="1.0" byte_order="LittleEndian">',
                 '  <Collection>']
        base = os.path.dirname(os.path.abspath(self.fname + ".pvd"))          # DataSet paths are relative to the .pvd
        lines += [f'    <DataSet timestep="{t!r}" part="0" file="{os.path.relpath(os.path.abspath(f), base)}"/>' for t, f in self.entries]
        lines += ['  </Collection>', '</VTKFile>', '']
        with open(self.fname + ".pvd", "w") as fh:
            fh.write("\n".join(lines))


def load_pvd(sim, fname="WaterLily.pvd", attrib=None, pressure="Pressure", velocity="Velocity"):
    """`load!(sim, Val(:pvd); fname, attrib)` (ext/WaterLilyReadVTKExt.jl:22-43): restart from the last file of the
    collection, set Δt[end] to the stored time (in flow units) and push CFL; returns a writer that appends."""
    base = os.path.dirname(os.path.abspath(fname))
    entries = [(t, os.path.relpath(f if os.path.isabs(f) else os.path.join(base, f))) for t, f in _read_pvd(fname)]
    t_last, f_last = entries[-1]
    if f_last.endswith(".pvti"):                           # written by a slab-decomposed run: reassemble, then take this rank's slab
        grid, data = read_pvti(f_last)
        n = getattr(sim, "nranks", 1)
        if n > 1:
            from . import slab_of
            data = {k: slab_of(a, sim.rank, n) for k, a in data.items()}
            grid = data[pressure].shape
    else:
        grid, data = read_vti(f_last)
    if tuple(grid) != tuple(sim.field_shape("p")):
        raise ValueError("The dimensions of the simulation do not match the dimensions of the vtk file.")
    sim.upload("p", data[pressure])
    sim.upload("u", data[velocity])
    dts = [np.float32(x) for x in sim.flow.Δt]
    dts[-1] = np.float32(t_last * sim.L / sim.U)
    dts.append(np.float32(sim.cfl()))
    sim.set_dts(dts)
    return VTKWriter(fname[:-4] if fname.endswith(".pvd") else fname, attrib, os.path.dirname(entries[0][1]) or ".",
                     entries, len(entries))
