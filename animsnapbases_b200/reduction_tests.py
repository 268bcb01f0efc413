"""Headless statement of the quantitative outputs of the reference's figure scripts:
generate_figures/pos_reduction_tests.py:22-96 and generate_figures/nl_reduction_tests.py:31-225.  Same CSV files (names,
headers, row order), no plotting (matplotlib / polyscope: unavailable on this box, and not needed for the numbers).
The reconstructions and their error norms stay on the device (engine.deim_reconstruct, engine.tensor_error_measures,
engine.reconstruction_errors); only the CSV rows reach the host.  Rank 0 writes the files.
"""
from __future__ import annotations

import csv
import os

import numpy as np

from . import runtime


def _rank0():
    _, comm = runtime.get()
    return comm.rank == 0


def pos_pca_tests(bases, param):
    """pos_reduction_tests.py:62-92: `<name>_posBases_xyz_fullBasesRange_K_singVals.csv` (row, x, y, z) with the normalised
    singular values of the bases per dimension, then the post-process self-checks.  Returns the (K, 3) array."""
    s = bases.test_basesSingVals()
    if _rank0():
        file_name = os.path.join(param.vertPos_output_directory, param.name + "_posBases_xyz_fullBasesRange_K_singVals")
        with open(file_name + '.csv', 'w', encoding='UTF8') as f:
            w = csv.writer(f)
            w.writerow(['row', 'x', 'y', 'z'])
            for row in range(s.shape[0]):
                w.writerow([row + 1, s[row, 0], s[row, 1], s[row, 2]])
    if param.q_orthogonal:
        bases.is_utmu_orthogonal()
    return s


def nl_pca_tests(bases):
    """nl_reduction_tests.py:75-90: `<name>_<constProj_name>_constrprojBases_xyz_fullBasesRange_Kp_singVals.csv`.  The
    reference writes the table twice, the second time as `[row, s[row, :]]` (an array repr in one cell); kept."""
    s = bases.test_basesSingVals()
    if _rank0():
        p = bases.param
        file_name = os.path.join(p.constProj_output_directory,
                                 p.name + "_" + p.constProj_name + "_constrprojBases_xyz_fullBasesRange_Kp_singVals")
        with open(file_name + '.csv', 'w', encoding='UTF8') as f:
            w = csv.writer(f)
            w.writerow(['i', 'x', 'y', 'z'])
            for row in range(s.shape[0]):
                w.writerow([row + 1, s[row, 0], s[row, 1], s[row, 2]])
            for row in range(s.shape[0]):
                w.writerow([row, s[row, :]])
    return s


def nl_postprocess_tests(bases):
    """nl_reduction_tests.py:108-115."""
    if bases.param.constProj_orthogonal:
        return bases.is_utmu_orthogonal()


def _case_file(bases, suffix):
    p = bases.param
    return os.path.join(p.constProj_output_directory, p.name + "_" + p.constProj_name + "_" +
                        p.constProj_bases_interpolation_type + "_" + p.constProj_basis_type + suffix)


def nl_reconstruction_test(bases, case, steps=1):
    """nl_reduction_tests.py:120-147: for r = 1, 1 + steps, ... <= numComp reconstruct the `case` ('train' / 'test') frames
    from r interpolation blocks and write `..._<case>_convergence_tests.csv` with
    numPoints, fro_error, max_err, relative_errors_x/y/z, relative3d.  Returns the rows."""
    rows = []
    for r in range(1, bases.numComp + 1, steps):
        print("Basis-blocks:", r)
        fro, mx, rel = bases.geom_reconstruction_errors(r, case)
        rows.append([r, fro, mx, rel[0], rel[1], rel[2], np.sum([rel[0], rel[1], rel[2]]) / 3])
    if _rank0():
        with open(_case_file(bases, "_%s_convergence_tests" % case) + '.csv', 'w', encoding='UTF8') as f:
            w = csv.writer(f)
            w.writerow(['numPoints', 'fro_error', 'max_err', 'relative_errors_x', 'relative_errors_y', 'relative_errors_z',
                        'relative3d'])
            w.writerows(rows)
    return rows


def nl_geom_tests(bases, steps=1):
    """nl_reduction_tests.py:188-225: train convergence, `..._num_interpol_elemnets.csv` (numPoints, num_elements), test
    convergence."""
    train = nl_reconstruction_test(bases, "train", steps)
    if _rank0():
        with open(_case_file(bases, "_num_interpol_elemnets") + '.csv', 'w', encoding='UTF8') as f:
            w = csv.writer(f)
            w.writerow(['numPoints', 'num_elements'])
            for r in range(1, bases.numComp + 1, steps):
                w.writerow([r, bases.geom_alpha_ranges[r - 1]])
    test = nl_reconstruction_test(bases, "test", steps)
    return train, test
