"""The host hand-over (staging, flattening, envelope: solver.cuh add_*_v, solver_finalize.cuh) runs on OpenMP threads. Its
result must not depend on how many: the bucketing of the incidence lists, the noise-group numbering, the PoseToPoint ranks
and the envelope minima are all defined by input order, not by which thread got there first. Same graph, 1 / 3 / all
threads: the per-factor linearisation is bit-identical, the LM trace and the values agree to the round-off of the device's
atomic sums (which vary from run to run on their own)."""
import ctypes
import os

import numpy as np
import pytest

import dynosam_b200 as d
import scenes

pytestmark = pytest.mark.gpu


def _solve(g, threads):
    ctypes.CDLL("libgomp.so.1").omp_set_num_threads(threads)
    p = d.default_lm_params()
    p.max_iterations = 3
    p.relative_error_tol = 0.0
    p.absolute_error_tol = 0.0
    lm = d.GpuLevenbergMarquardt(None, p)
    lm.load_host_arrays(g)
    e0 = lm.error()
    lin = lm.linearize()
    pose, pts = lm.optimize()
    return e0, list(lm.error_trace), pose, pts, lm.timing(), lin


@pytest.mark.parametrize("scene", ["small_noisy", "medium_noisy"])
def test_hand_over_is_independent_of_the_thread_count(scene):
    g = getattr(scenes, scene)().full_batch_graph()
    all_threads = max(2, os.cpu_count() or 2)
    try:
        ref = _solve(g, 1)
        for n in (3, all_threads):
            got = _solve(g, n)
            for k in ref[5]:
                assert np.array_equal(got[5][k], ref[5][k]), k
            assert got[0] == pytest.approx(ref[0], rel=1e-12)
            np.testing.assert_allclose(got[1], ref[1], rtol=1e-9)
            np.testing.assert_allclose(got[2], ref[2], atol=1e-8)
            np.testing.assert_allclose(got[3], ref[3], atol=1e-8)
            assert got[4]["half_bandwidth"] == ref[4]["half_bandwidth"] and got[4]["reduced_dim"] == ref[4]["reduced_dim"]
    finally:
        ctypes.CDLL("libgomp.so.1").omp_set_num_threads(all_threads)
