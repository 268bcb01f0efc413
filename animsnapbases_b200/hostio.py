"""Host-side file formats and bookkeeping of the basis pipeline (the reference's `utils/utils.py` surface that
`posComponents` / `constraintsComponents` use).  Only byte layouts and messages are reproduced; the per-element
`struct.pack` loops of the reference (0.68 us / element) become one contiguous `tofile`.

  store_components            utils/utils.py:14-38      `.bin`: <i N><i 3K> + f64 LE in order d, k, i;  `.npy`: np.save
  store_interpol_points_vector utils/utils.py:77-98     `.bin`: <i n><i 1> + indices as f64
  read_mass_bin               snapbases/posSnapshots.py:142-149, snapbases/nonlinear_snapshots.py:181-192
  log_time                    utils/utils.py:209-237    function_timings.txt (first write of the process truncates)
  read_animation / write_components_animation           the `.h5` containers of posSnapshots.read (:108-121) and
      posComponents.store_animations (:330-341); h5py/libhdf5 are not in this image, so an `.npz` with the SAME keys is
      read / written when h5py cannot be imported.
"""
from __future__ import annotations

import functools
import os
import struct
import time

import numpy as np

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


def _h5py():
    try:
        import h5py
        return h5py
    except Exception:
        return None


def _npz_sidecar(path):
    return path if path.endswith(".npz") else os.path.splitext(path)[0] + ".npz"


def read_animation(path):
    """-> (verts (F, N, 3) in the stored dtype, tris) from `<path>` (.h5 with h5py) or its `.npz` twin."""
    h5 = _h5py()
    if h5 is not None and os.path.exists(path) and not path.endswith(".npz"):
        with h5.File(path, "r") as f:
            return f["verts"][()], f["tris"][()]
    data = np.load(_npz_sidecar(path))
    return data["verts"], data["tris"]


def write_animation(path, verts, tris):
    h5 = _h5py()
    if h5 is not None and not path.endswith(".npz"):
        with h5.File(path, "w") as f:
            f.create_dataset("verts", data=np.asarray(verts, dtype=np.float32), compression="gzip")
            f["tris"] = tris
        return path
    out = _npz_sidecar(path)
    np.savez(out, verts=np.asarray(verts, dtype=np.float32), tris=tris)
    return out


def write_components_animation(path, default_verts, tris, comps):
    """components.h5 of posComponents.store_animations: keys `default`, `tris`, `comp%03d`."""
    h5 = _h5py()
    if h5 is not None and not path.endswith(".npz"):
        with h5.File(path, "w") as f:
            f["default"] = default_verts
            f["tris"] = tris
            for i, c in enumerate(comps):
                f["comp%03d" % i] = c
        return path
    out = _npz_sidecar(path)
    np.savez(out, default=default_verts, tris=tris, **{"comp%03d" % i: c for i, c in enumerate(comps)})
    return out
