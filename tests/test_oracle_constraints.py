"""CPU: the oracle's constraint / add_constant_value restatement pinned to the
outputs of the reference's own source (tests/golden/make_golden_constraints.py)
and to the reference's known answer tests/api/test_features.py:436-455."""
import os

import numpy as np
import pytest

from oracle import ocsmesh_oracle as O

GOLD = os.path.join(os.path.dirname(__file__), "golden",
                    "reference_golden_constraints.npz")

OPS = {
    "constant": lambda h: h.add_constant_value(123.456, -3.0, 2.5),
    "topo_min_rate": lambda h: h.add_topo_bound_constraint(
        400.0, upper_bound=3.0, lower_bound=-8.0, value_type="min", rate=0.02),
    "topo_max_norate": lambda h: h.add_topo_bound_constraint(
        90.0, upper_bound=3.0, lower_bound=-8.0, value_type="max", rate=None),
    "func_min_norate": lambda h: h.add_topo_func_constraint(
        upper_bound=-1.0, value_type="min", rate=None),
    "courant": lambda h: h.add_courant_num_constraint(
        upper_bound=0.8, lower_bound=0.3, timestep=20, wave_amplitude=1.5),
}


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def case_names(gold):
    return sorted(k[:-5] for k in gold.files if k.endswith("_args"))


def opt_int(v):
    return None if np.isnan(v) else int(v)


def lines_of(gold):
    return [gold[f"feat_line_{k}"] for k in range(int(gold["feat_n_lines"]))]


def test_apply_rate_known_answer(gold):
    old = np.array([0, 0, 0, 0, 10, 10, 10, 10, 20, 20, 20, 20])
    loc = np.array([(i, 0) for i in range(len(old))])
    mask = np.zeros(len(old), bool)
    mask[3] = True
    v = old.copy()
    v[mask] = 5
    got = O.apply_rate(v, loc, mask, 0.199, -1, np.greater)
    check = old.copy()
    check[:4] = [2, 3, 4, 5]          # tests/api/test_features.py:452-453
    np.testing.assert_array_equal(got, check)
    np.testing.assert_array_equal(got, gold["rate_known"])


def test_oracle_matches_reference_source(gold):
    names = case_names(gold)
    assert len(names) >= 2
    for name in names:
        chunk, ovl = (opt_int(v) for v in gold[f"{name}_args"])

        def fresh():
            r = O.OracleRaster(gold["dem"], gold["dem_transform"],
                               chunk_size=chunk, overlap=ovl)
            oh = O.OracleHfun(r, 40.0, 2000.0)
            oh.add_feature(lines_of(gold), 0.01, 60.0)
            return oh
        oh = fresh()
        np.testing.assert_array_equal(oh.values, gold[f"{name}_seed"])
        for op, fn in OPS.items():
            oh = fresh()
            fn(oh)
            np.testing.assert_array_equal(oh.values, gold[f"{name}_{op}"],
                                          err_msg=f"{name}/{op}")
            if op == "topo_min_rate":
                oh.add_subtidal_flow_limiter(30.0, 900.0, None, None)
                np.testing.assert_array_equal(
                    oh.values, gold[f"{name}_topo_min_rate_then_limiter"])


def test_msh_t_sampling_matches_reference_source(gold):
    """utils.interpolate(method='nearest') lifted from ocsmesh/utils.py:1122-1140."""
    col, row, w, h = (int(v) for v in gold["msh_win"])
    r = O.OracleRaster(gold["dem"], gold["dem_transform"])
    win = O.Window(col, row, w, h)
    vals = gold["msh_hfun"][row:row + h, col:col + w]
    for step in (1, 3):
        c, v = O.msh_t_source(vals, r.get_xy(win), h, w, 80.0, 900.0, step)
        got = O.interpolate_nearest(c, v, gold["msh_nodes"])
        np.testing.assert_array_equal(got, gold[f"msh_nearest_stride{step}"])
