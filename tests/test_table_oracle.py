"""CPU checks of the table-side restatement and of the host logic around the table kernels."""
import ctypes as C

import numpy as np

from irbfn_b200 import _lib
from oracle import table_oracle as tor


def test_remove_infeasible_keeps_rows_with_any_valid_output():
    x = np.arange(12.0).reshape(4, 3)
    y = np.array([[1.0, 2.0], [-999.0, -999.0], [-999.0, 5.0], [3.0, 4.0]])
    xv, yv = tor.remove_infeasible(x, y)
    assert xv.tolist() == x[[0, 2, 3]].tolist() and yv.tolist() == y[[0, 2, 3]].tolist()


def test_epoch_batches_drop_the_ragged_tail():
    b = tor.epoch_batches(np.arange(10)[::-1], 10, 4)
    assert b.tolist() == [[9, 8, 7, 6], [5, 4, 3, 2]]


def test_nearest_row_and_grid_index_small_cases():
    table = np.array([[0.0, 0.0], [1.0, 0.0], [0.0, 0.0]])
    idx, d1, d2 = tor.nearest_row(np.array([[0.1, 0.0], [0.9, 0.1]]), table)
    assert idx.tolist() == [0, 1] and np.isclose(d1[0], 0.1) and np.isclose(d2[0], 0.1)
    g = tor.grid_index(np.array([[-1.0, 0.5], [0.0, 2.0], [9.0, 1.0]]), [[0.0, 1.0, 2.0], [1.0, 2.0]])
    assert g.tolist() == [[0, 0], [1, 1], [2, 1]]


def test_table_entry_points_validate_before_touching_a_device():
    lib = _lib.load()
    one = C.c_void_p(16)
    assert lib.irbfn_gather_rows(one, one, one, one, one, 4, 0, 8, 2, one, None) == -1
    assert lib.irbfn_gather_rows(one, one, one, one, one, 0, 10, 8, 2, one, None) == 0
    assert lib.irbfn_gather_rows(one, None, one, one, one, 4, 10, 8, 2, one, None) == -4
    assert lib.irbfn_grid_index(one, one, one, one, 4, 0, None) == -1
    assert lib.irbfn_grid_index_f64(one, one, one, one, 4, 0, None) == -1
    assert lib.irbfn_grid_index_f64(one, None, one, one, 4, 2, None) == -4
    assert lib.irbfn_grid_index_f64(one, one, one, one, 0, 2, None) == 0
    assert lib.irbfn_nearest_row(one, one, one, None, 4, 0, 8, None, 0, None) == -1
    assert lib.irbfn_nearest_row(one, one, one, None, 4, 10, 9, None, 0, None) == -2
    assert lib.irbfn_nearest_row(one, one, one, None, 0, 10, 8, None, 0, None) == 0
    # small batch: the table is split over the 148 SMs -> 148 partial (index, distance) rows
    assert lib.irbfn_nearest_row_workspace_bytes(400, 1 << 20) == 148 * 400 * 12
    assert lib.irbfn_nearest_row_workspace_bytes(1 << 20, 1 << 20) == (1 << 20) * 12
    assert lib.irbfn_jacobian(one, one, one, one, None, None, None, one, None, None, 4, 2, 3, 4, 2, 0, None) == -5
    assert lib.irbfn_jacobian(one, one, one, one, None, one, None, one, None, None, 4, 2, 3, 9, 2, 0, None) == -2
    fl = (C.c_float * 13)()
    assert lib.irbfn_dynamics_loss(9, one, one, one, fl, one, one, 4, 8, 2, 4, one, 1 << 20, None) == -1
    assert lib.irbfn_dynamics_loss(1, one, one, one, fl, one, one, 4, 8, 2, 4, one, 1 << 20, None) == -1
    assert lib.irbfn_integrate(0, one, fl, one, 4, 17, None) == -1
