"""The oracle (and the CUDA path) against GTSAM itself on the reference's hand-built graph -- when
tests/golden/three_frame_gtsam.json exists. That file is written by tools/gtsam_pin/pin_three_frame.cc on a machine that has
GTSAM 4.2 and the reference (this image has neither); until someone commits it these tests skip, and DESIGN.md keeps saying
"parity unpinned against GTSAM"."""
import json
import os

import numpy as np
import pytest

import dynosam_b200 as d
import scenes
from oracle_lib import OracleGraph

PIN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "three_frame_gtsam.json")
needs_pin = pytest.mark.skipif(not os.path.exists(PIN), reason="tests/golden/three_frame_gtsam.json not generated: needs GTSAM 4.2 + the reference (tools/gtsam_pin/README.md)")


def _pin(which):
    with open(PIN) as f:
        return json.load(f)[which]


def _check(result, pin):
    assert (result["iterations"], result["inner_iterations"]) == (pin["iterations"], pin["inner_iterations"])
    assert np.allclose(result["error_trace"], pin["error_trace"], rtol=1e-6, atol=1e-12)
    assert np.allclose(result["pose_Rt"], np.asarray(pin["pose_Rt"]), rtol=0, atol=1e-6)
    assert np.allclose(result["point_xyz"], np.asarray(pin["point_xyz"]), rtol=0, atol=1e-6)


@needs_pin
@pytest.mark.parametrize("which", ["graph", "full_graph"])
def test_oracle_matches_gtsam(which):
    g, pin = scenes.three_frame(which), _pin(which)
    ora = OracleGraph.from_graph(g)
    assert abs(ora.error() - pin["initial_error"]) <= 1e-9 * pin["initial_error"]
    r = ora.optimize()
    r["pose_Rt"], r["point_xyz"] = ora.values()
    _check(r, pin)


@needs_pin
@pytest.mark.gpu
@pytest.mark.parametrize("which", ["graph", "full_graph"])
def test_gpu_matches_gtsam(which):
    g, pin = scenes.three_frame(which), _pin(which)
    lm = d.GpuLevenbergMarquardt(g, d.default_lm_params())
    assert abs(lm.error() - pin["initial_error"]) <= 1e-9 * pin["initial_error"]
    pose, point = lm.optimize()
    _check({"iterations": lm.iterations(), "inner_iterations": lm.getInnerIterations(), "error_trace": lm.error_trace, "pose_Rt": pose, "point_xyz": point}, pin)


def test_pin_program_reads_the_fixture_fields_that_exist():
    """The C++ program cannot be compiled here; at least the fixture fields it reads must exist under the names it uses."""
    src = open(os.path.join(os.path.dirname(PIN), "..", "..", "tools", "gtsam_pin", "pin_three_frame.cc")).read()
    with open(os.path.join(os.path.dirname(PIN), "three_frame.json")) as f:
        fx = json.load(f)
    for which in ("graph", "full_graph"):
        g = fx[which]
        for key in ("pose_keys", "pose_Rt", "point_keys", "point_xyz", "point_sigma", "pose_sigma", "p2p", "ternary", "between", "prior"):
            assert key in g and ('"%s"' % key) in src
        for lst, fields in (("p2p", ("order", "pose", "point", "z")), ("ternary", ("order", "prev", "cur", "motion")), ("between", ("order", "a", "b", "z")), ("prior", ("order", "a", "z"))):
            assert all(set(fields) <= set(f) for f in g[lst])
        orders = sorted(f["order"] for lst in ("p2p", "ternary", "between", "prior") for f in g[lst])
        assert orders == list(range(len(orders)))
