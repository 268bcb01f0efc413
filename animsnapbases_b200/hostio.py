"""Host-side file formats and bookkeeping of the basis pipeline (the reference's `utils/utils.py` surface that
`posComponents` / `constraintsComponents` use).  Only byte layouts and messages are reproduced; the per-element
`struct.pack` loops of the reference (0.68 us / element) become one contiguous `tofile`.

  store_components            utils/utils.py:14-38      `.bin`: <i N><i 3K> + f64 LE in order d, k, i;  `.npy`: np.save
  store_interpol_points_vector utils/utils.py:77-98     `.bin`: <i n><i 1> + indices as f64
  read_mass_bin               snapbases/posSnapshots.py:142-149, snapbases/nonlinear_snapshots.py:181-192
  log_time                    utils/utils.py:209-237    function_timings.txt (first write of the process truncates)
  read_animation / write_components_animation           the `.h5` containers of posSnapshots.read (:108-121) and
      posComponents.store_animations (:330-341): real HDF5 files through h5mini.py (h5py / libhdf5 are not in this image,
      SURVEY H8); a name ending in `.npz` selects the numpy container with the same keys.
"""
from __future__ import annotations

import functools
import os
import struct
import time

import numpy as np

from . import h5mini

_has_written = False


def log_time(file_path=""):
    """Decorator with the reference's output format: prints and appends `Function '<name>' executed in <t> seconds.`"""

    def decorator(func):
        @functools.wraps(func)
        def wrapper(*args, **kwargs):
            global _has_written
            mode = "a" if _has_written else "w"
            _has_written = True
            t0 = time.time()
            result = func(*args, **kwargs)
            dt = time.time() - t0
            print(f"Function '{func.__name__}' executed in {dt:.4f} seconds.")
            with open(file_path + "function_timings.txt", mode) as f:
                f.write(f"Function '{func.__name__}' executed in {dt:.4f} seconds.\n")
            return result

        return wrapper

    return decorator


def components_bin_name(file_name, F, K, extension=".bin", col_name="K"):
    if extension == ".bin":
        return file_name + "F" + str(F) + col_name + str(K) + extension
    return file_name + str(F) + "K" + str(K) + ".npy"          # the reference drops the 'F' for .npy (utils/utils.py:38)


def store_components_packed(file_name, F, K, N, dim, packed_dkn, extension=".bin", col_name="K"):
    """`packed_dkn`: host array (dim, Kfull, N) = basesTensor.transpose(2, 0, 1) (asb_pack_bin); writes components [0, K)."""
    assert packed_dkn.shape[0] == dim and packed_dkn.shape[1] >= K and packed_dkn.shape[2] == N
    if extension == ".bin":
        with open(components_bin_name(file_name, F, K, extension, col_name), "wb") as f:
            f.write(struct.pack("<i", N))
            f.write(struct.pack("<i", dim * K))
            for d in range(dim):
                np.ascontiguousarray(packed_dkn[d, :K], dtype="<f8").tofile(f)
    if extension == ".npy":
        np.save(file_name + str(F) + "K" + str(K), np.ascontiguousarray(packed_dkn[:, :K].transpose(1, 2, 0)))


def store_components(file_name, F, K, N, dim, bases_tensor, extension=".bin", col_name="K"):
    """Same signature as the reference (utils/utils.py:14): bases_tensor (K, N, dim) host array."""
    assert bases_tensor.shape == (K, N, dim)
    store_components_packed(file_name, F, K, N, dim, np.ascontiguousarray(bases_tensor.transpose(2, 0, 1)), extension, col_name)


def store_interpol_points_vector(file_name, F, K, points, extension=".bin", col_name="K"):
    points = np.asarray(points)
    assert K <= points.shape[0]
    print("Storing ", points.shape[0], "interpolation points")
    if extension == ".bin":
        with open(file_name + "F" + str(F) + col_name + str(K) + "_points" + str(points.shape[0]) + extension, "wb") as f:
            f.write(struct.pack("<i", points.shape[0]))
            f.write(struct.pack("<i", 1))
            points.astype("<f8").tofile(f)
    if extension == ".npy":
        np.save(file_name + str(F) + "K" + str(K) + "_points" + str(points.shape[0]), points)


def load_vector_values(full_filename):
    with open(full_filename, "rb") as f:
        n = struct.unpack("<i", f.read(4))[0]
        f.read(4)
        return np.fromfile(f, dtype="<f8", count=n)


def read_mass_bin(path, expected=None):
    """<i N><i 1> + N f64 LE."""
    with open(path, "rb") as f:
        n = struct.unpack("<i", f.read(4))[0]
        f.read(4)
        if expected is not None:
            assert n == expected
        return np.fromfile(f, dtype="<f8", count=n).astype(np.float64)


def write_mass_bin(path, mass):
    with open(path, "wb") as f:
        f.write(struct.pack("<i", len(mass)))
        f.write(struct.pack("<i", 1))
        np.asarray(mass, dtype="<f8").tofile(f)


def _npz_sidecar(path):
    return path if path.endswith(".npz") else os.path.splitext(path)[0] + ".npz"


class MemoryAnimation:
    """An animation handed over in host memory instead of a file (registered under a path with `register_animation`, so the
    reference's config object keeps naming its snapshot 'files' as before).  `verts`: (F, n_local, 3) float32 / float64, a
    numpy array or a (pinned) torch tensor holding the rows [row0, row0 + n_local) of the N-vertex animation -- a rank of a
    row-sharded run registers only its shard; `rest`: the complete frame 0 (N, 3), which the geodesic set-up needs on every
    rank; `tris` (T, 3)."""

    def __init__(self, verts, tris, rest=None, row0=0, n_global=None):
        self.verts, self.tris, self.row0 = verts, np.asarray(tris), int(row0)
        self.frames, self.n_local = int(verts.shape[0]), int(verts.shape[1])
        self.n_global = self.n_local if n_global is None else int(n_global)
        self.rest = np.asarray(verts[0], dtype=np.float64) if rest is None else np.asarray(rest, dtype=np.float64)
        assert self.rest.shape == (self.n_global, 3), "rest must be the complete frame 0"
        self.shape = (self.frames, self.n_global, 3)

    def rows(self, row0, n):
        """The stored rows [row0, row0 + n) -- must lie inside what this rank registered."""
        a = row0 - self.row0
        assert 0 <= a and a + n <= self.n_local, "this rank did not register the rows of its shard"
        return self.verts[:, a:a + n]

    def complete(self):
        assert self.n_local == self.n_global, "only a shard of the animation is in this rank's memory"
        v = self.verts
        return v.numpy() if hasattr(v, "numpy") else np.asarray(v)


_MEMORY = {}


def register_animation(path, animation):
    """Make `read_animation(path)` return the MemoryAnimation (None removes the entry)."""
    if animation is None:
        _MEMORY.pop(path, None)
    else:
        _MEMORY[path] = animation


def read_animation(path):
    """-> (verts (F, N, 3) in the stored dtype, tris) from `<path>` (.h5 with h5py) or its `.npz` twin; a path registered
    with `register_animation` yields its MemoryAnimation instead of an array."""
    if path in _MEMORY:
        return _MEMORY[path], _MEMORY[path].tris
    if not path.endswith(".npz") and h5mini.is_hdf5(path):     # a real HDF5 container (h5py's gzip-chunked `verts` included)
        d = h5mini.read_datasets(path, ("verts", "tris"))
        return d["verts"], d["tris"]
    data = np.load(_npz_sidecar(path))
    return data["verts"], data["tris"]


def write_animation(path, verts, tris, attrs=None):
    """`verts` (float32) / `tris` to `path`: an HDF5 file (h5mini; contiguous datasets) unless the name ends in .npz."""
    if not path.endswith(".npz"):
        return h5mini.write_datasets(path, {"verts": np.asarray(verts, dtype=np.float32), "tris": np.asarray(tris)}, attrs)
    out = _npz_sidecar(path)
    np.savez(out, verts=np.asarray(verts, dtype=np.float32), tris=tris)
    return out


def write_components_animation(path, default_verts, tris, comps):
    """components.h5 of posComponents.store_animations: keys `default`, `tris`, `comp%03d`."""
    data = {"default": np.asarray(default_verts), "tris": np.asarray(tris)}
    data.update({"comp%03d" % i: np.asarray(c) for i, c in enumerate(comps)})
    if not path.endswith(".npz"):
        return h5mini.write_datasets(path, data)           # what utils/process.py:367-376 (load_splocs) and the viewers read
    out = _npz_sidecar(path)
    np.savez(out, **data)
    return out


# ------------------------------------------------------------------------------------------------ sparse operators (n3)
_TRIPLET = np.dtype([("row", "<i4"), ("col", "<i4"), ("val", "<f8")])


def read_sparse_matrix_from_bin(filename):
    """`<i rows><i cols><i nnz>` + nnz x (`<i row><i col><d value>`) -> scipy CSR, duplicates summed (utils/utils.py:289-312,
    which unpacks the 3 nnz fields one struct at a time; here one structured read)."""
    from scipy.sparse import csr_matrix
    with open(filename, "rb") as f:
        rows, cols, nnz = struct.unpack("<iii", f.read(12))
        t = np.fromfile(f, dtype=_TRIPLET, count=nnz)
    assert len(t) == nnz, "truncated sparse matrix file"
    return csr_matrix((t["val"], (t["row"], t["col"])), shape=(rows, cols))


def write_sparse_matrix_to_bin(filename, A):
    """Inverse of read_sparse_matrix_from_bin (the reference's files come from its C++ side; this one serves tests and
    synthetic configs)."""
    A = A.tocoo()
    t = np.empty(A.nnz, dtype=_TRIPLET)
    t["row"], t["col"], t["val"] = A.row, A.col, A.data
    with open(filename, "wb") as f:
        f.write(struct.pack("<iii", A.shape[0], A.shape[1], A.nnz))
        t.tofile(f)


def read_sparse_matrix(file_name, type, key=None):
    """utils/utils.py:314-323: '.bin' triplet file, or '.npz' holding a pickled scipy sparse matrix under `key` (returned
    as the 0-d object array np.load yields, like the reference; constraintsComponents unwraps it)."""
    if type == ".bin":
        return read_sparse_matrix_from_bin(file_name)
    elif type == ".npz":
        if key is None:
            raise ValueError("Empty file or wrong key:", type)
        return np.load(file_name, allow_pickle=True)[key]
    raise ValueError("unknown sparse matrix file_type:", type)


# ------------------------------------------------------------------------------------------------ snapshot ingest (a1, n3)
def load_off(filename, no_colors=False):
    """OFF / COFF reader of utils/process.py:46-66: (verts (n, 3) float64, [colors uint8,] faces (m, k) int or None)."""
    from io import StringIO
    lines = [ln for ln in open(filename).readlines() if ln.strip() != "" and ln[0] != "#"]
    assert lines[0].strip() in ["OFF", "COFF"], "OFF header missing"
    has_colors = lines[0].strip() == "COFF"
    n_verts, n_faces, _ = map(int, lines[1].split())
    vertex_data = np.loadtxt(StringIO("".join(lines[2:2 + n_verts])), dtype=float, ndmin=2)
    faces = np.loadtxt(StringIO("".join(lines[2 + n_verts:])), dtype=int, ndmin=2)[:, 1:] if n_faces > 0 else None
    colors = None
    if has_colors:
        colors = vertex_data[:, 3:].astype(np.uint8)
        vertex_data = vertex_data[:, :3]
    return (vertex_data, faces) if no_colors else (vertex_data, colors, faces)


_PLY_TYPES = {"char": "i1", "uchar": "u1", "short": "i2", "ushort": "u2", "int": "i4", "uint": "u4", "float": "f4", "double": "f8",
              "int8": "i1", "uint8": "u1", "int16": "i2", "uint16": "u2", "int32": "i4", "uint32": "u4", "float32": "f4", "float64": "f8"}


def load_ply(filename):
    """-> (verts (N, 3) float, tris (T, 3) int) of a triangle-mesh PLY file, ascii or binary (utils/process.py:30-44 reads it
    through TVTK's PLYReader, a third-party dependency that is absent here; this reader needs none).  Vertex properties
    other than x, y, z are skipped; every face must be a triangle, as the reference asserts (:43)."""
    with open(filename, "rb") as f:
        assert f.readline().strip() == b"ply", "PLY header missing"
        fmt, elements = None, []
        while True:
            tok = f.readline().decode("ascii", "replace").split()
            if not tok or tok[0] == "comment" or tok[0] == "obj_info":
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                elements.append((tok[1], int(tok[2]), []))
            elif tok[0] == "property":
                elements[-1][2].append(tok[1:])
            elif tok[0] == "end_header":
                break
        assert fmt in ("ascii", "binary_little_endian", "binary_big_endian"), "unknown PLY format"
        end = ">" if fmt == "binary_big_endian" else "<"
        verts = tris = None
        for name, count, props in elements:
            is_list = any(p[0] == "list" for p in props)
            if fmt == "ascii":
                rows = [f.readline().split() for _ in range(count)]
                if name == "vertex":
                    names = [p[-1] for p in props]
                    cols = [names.index(c) for c in ("x", "y", "z")]
                    verts = np.array([[float(r[c]) for c in cols] for r in rows], dtype=np.float64)
                elif name == "face":
                    assert all(int(r[0]) == 3 for r in rows), "only triangle meshes"
                    tris = np.array([[int(v) for v in r[1:4]] for r in rows], dtype=np.int64)
            elif not is_list:
                dt = np.dtype([(p[-1], end + _PLY_TYPES[p[0]]) for p in props])
                data = np.frombuffer(f.read(dt.itemsize * count), dtype=dt)
                if name == "vertex":
                    verts = np.stack([data["x"], data["y"], data["z"]], axis=1).astype(np.float64)
            else:
                assert len(props) == 1 or name != "face", "faces with extra properties are not supported"
                _, ct, it, _ = props[0]
                dt = np.dtype([("n", end + _PLY_TYPES[ct]), ("v", end + _PLY_TYPES[it], (3,))])
                data = np.frombuffer(f.read(dt.itemsize * count), dtype=dt)
                assert np.all(data["n"] == 3), "only triangle meshes"
                if name == "face":
                    tris = data["v"].astype(np.int64)
    assert verts is not None and tris is not None, "PLY file without vertex / face elements"
    return verts, tris


def filter_reindex(condition, target):
    """utils/process.py:94-106: indices `target` re-numbered after dropping the entries where `condition` is False."""
    if condition.dtype != bool:
        raise ValueError("condition must be a binary array")
    return (np.cumsum(condition) - 1)[target]


def preprocess_mesh_animation(verts, tris):
    """utils/process.py:107-148: drop zero-area triangles (frame 0, |n| <= 1e-8), keep the largest connected component,
    subtract the global mean and divide by the largest per-frame extent (-0.5 .. 0.5 cube; sets the scale of the support-map
    distances).  The reference's `verts.ptp(axis=1)` (removed in numpy 2) is `max - min` over the vertices of each frame.
    verts (F, N, 3) float, tris (T, 3) int -> (verts', tris', dropped-vertex mask, mean (3,), scale).  Host, once per dataset."""
    from scipy.sparse import csr_matrix
    from scipy.sparse.csgraph import connected_components
    verts = np.array(verts, copy=True)
    assert verts.ndim == 3 and tris.ndim == 2
    e1 = verts[0, tris[:, 1]] - verts[0, tris[:, 0]]
    e2 = verts[0, tris[:, 2]] - verts[0, tris[:, 0]]
    tris = tris[np.linalg.norm(np.cross(e1, e2), axis=1) > 1.e-8]
    ij = np.r_[np.c_[tris[:, 0], tris[:, 1]], np.c_[tris[:, 0], tris[:, 2]], np.c_[tris[:, 1], tris[:, 2]]]
    G = csr_matrix((np.ones(len(ij)), ij.T), shape=(verts.shape[1], verts.shape[1]))
    n_components, labels = connected_components(G, directed=False)
    if n_components > 1:
        size_components = np.bincount(labels)
        print("[warning] found %d connected components in the mesh, keeping only the biggest one" % n_components)
        print("component sizes: ")
        print(size_components)
        keep_vert = labels == size_components.argmax()
    else:
        keep_vert = np.ones(verts.shape[1], bool)
    verts = verts[:, keep_vert, :]
    tris = filter_reindex(keep_vert, tris[keep_vert[tris].all(axis=1)])
    verts_mean = verts.mean(axis=0).mean(axis=0)
    verts -= verts_mean
    verts_scale = np.abs(verts.max(axis=1) - verts.min(axis=1)).max()
    verts /= verts_scale
    return verts, tris, ~keep_vert, verts_mean, verts_scale


def convert_sequence_to_hdf5(filename_pattern, loader_function, hdf_output_file, max_frames, icreament):
    """utils/process.py:68-92: a numbered sequence of mesh files -> one animation file (`verts` float32, `tris`, attributes
    `mean`, `scale`).  Written as .h5 when h5py is present, otherwise as the `.npz` twin `read_animation` understands."""
    import re
    from glob import glob
    files = glob(os.path.expanduser(filename_pattern))
    files.sort(key=lambda s: [int(t) if t.isdigit() else t for t in re.split(r"([0-9]+)", s)])    # sort_nicely
    verts_all, tris, count = [], None, 0
    for i, f in enumerate(files):
        if i % icreament == 0 and count < max_frames:
            print("loading file %d/%d [%s]" % (i + 1, len(files), f))
            verts, new_tris = loader_function(f)
            if tris is not None and new_tris.shape != tris.shape:
                raise ValueError("inconsistent topology between meshes of different frames")
            tris = new_tris
            verts_all.append(verts)
            count += 1
    verts_all = np.array(verts_all, np.float32)
    verts_all, tris, _, verts_mean, verts_scale = preprocess_mesh_animation(verts_all, tris)
    if not hdf_output_file.endswith(".npz"):
        out = write_animation(hdf_output_file, verts_all, tris, attrs={"mean": verts_mean, "scale": np.float64(verts_scale)})
    else:
        out = _npz_sidecar(hdf_output_file)
        np.savez(out, verts=verts_all, tris=tris, mean=verts_mean, scale=verts_scale)
    print("saved as %s" % out)
    return out


def align(input_hdf5_file, output_hdf5_file, rigid):
    """utils/process.py:235-250: Procrustes-align every frame of an animation file to frame 0 (rigid: rotation + translation,
    otherwise translation only) and write `verts` (float32) / `tris` to the output file.  The per-frame centroid / 3x3
    covariance reductions and the transform pass run on the device (asb_align_centroids / _cov / _solve / _apply); the
    reference's Python loop over frames with a 4x4 homogeneous matmul each is gone."""
    from . import runtime
    ops, _ = runtime.get()
    verts, tris = read_animation(input_hdf5_file)
    out, _ = ops.align(ops.to_device(np.ascontiguousarray(verts, dtype=np.float32)), rigid)
    return write_animation(output_hdf5_file, out.cpu().numpy(), tris)
