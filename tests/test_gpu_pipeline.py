"""PipelinedLevenbergMarquardt: a stream of graphs solved double buffered (the hand-over of graph k+1 beside the solve of
graph k, two handles, one host thread) gives, in order, exactly what solving them one at a time gives."""
import numpy as np
import pytest

import dynosam_b200 as d

pytestmark = pytest.mark.gpu


def test_pipelined_stream_equals_one_at_a_time():
    scene = d.Scene.synthetic(n_frames=60, n_objects=3, static_per_frame=60, dynamic_per_object=12, object_span=30, seed=9)
    graphs = [scene.window_graph(a, a + 20, True) for a in (0, 8, 16, 24, 32)] + [scene.full_batch_graph()]
    p = d.default_lm_params()
    serial = []
    lm = d.GpuLevenbergMarquardt(None, p)
    for g in graphs:
        lm.load_host_arrays(g)
        pose, point = lm.optimize()
        serial.append((pose, point, lm.iterations(), list(lm.error_trace)))
    pipe = d.PipelinedLevenbergMarquardt(p)
    out = list(pipe.solve_all(graphs))
    assert len(out) == len(graphs)
    for (pose, point, its, trace), (pose0, point0, its0, trace0) in zip(out, serial):
        assert its == its0 and trace == trace0
        assert np.array_equal(pose, pose0) and np.array_equal(point, point0)
    # a second pass through the same object, by graph handle this time (sums may differ in the last bit: the pose-only
    # factors arrive grouped by noise model through the host arrays), an empty stream
    again = list(pipe.solve_all(iter(graphs[:2]), graph_handles=True))
    assert again[1][2] == serial[1][2] and np.allclose(again[1][0], serial[1][0], rtol=0, atol=1e-9)
    assert list(pipe.solve_all([])) == []
    pipe.close()
