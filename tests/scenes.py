"""Seeded scenes shared by the parity tests (sizes the CPU oracle finishes in seconds)."""
import dynosam_b200 as d


def simulator_cfg1(n_frames=50):
    """BASELINE config 1: the reference's makeDefaultScenario(), 2 objects, 50 frames."""
    return d.Scene.simulator(n_frames)


def small_noisy(seed=7, n_frames=40, n_objects=3, static_per_frame=50, dynamic_per_object=12):
    return d.Scene.synthetic(n_frames=n_frames, n_objects=n_objects, static_per_frame=static_per_frame,
                             dynamic_per_object=dynamic_per_object, static_track_max=20, dynamic_track_max=15, seed=seed)


def medium_noisy(seed=11):
    return d.Scene.synthetic(n_frames=120, n_objects=4, static_per_frame=120, dynamic_per_object=25, seed=seed)


def three_frame(which="graph"):
    """The reference's hand-built graph (test/test_rgbd_backend.cc:272-664) from tests/golden/three_frame.json as the
    flat arrays every consumer here takes (GpuLevenbergMarquardt.load_host_arrays, OracleGraph.from_graph, np_ref)."""
    import json
    import os
    from types import SimpleNamespace
    import numpy as np
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "three_frame.json")) as f:
        t = json.load(f)[which]
    g = SimpleNamespace()
    g.pose_keys = np.array([int(k) for k in t["pose_keys"]], dtype=np.uint64)
    g.point_keys = np.array([int(k) for k in t["point_keys"]], dtype=np.uint64)
    g.pose_Rt, g.point_xyz = np.array(t["pose_Rt"], dtype=np.float64), np.array(t["point_xyz"], dtype=np.float64)
    g.n_pose, g.n_point = len(g.pose_keys), len(g.point_keys)
    i32 = lambda rows, k: np.array([r[k] for r in rows], dtype=np.int32)   # noqa: E731
    f64 = lambda rows, k, w: np.array([r[k] for r in rows], dtype=np.float64).reshape(len(rows), w)   # noqa: E731
    p, q, b, pr = t["p2p"], t["ternary"], t["between"], t["prior"]
    g.p2p_pose, g.p2p_point, g.p2p_order, g.p2p_z = i32(p, "pose"), i32(p, "point"), i32(p, "order"), f64(p, "z", 3)
    g.p2p_sigma, g.p2p_huber = np.full(len(p), t["point_sigma"]), np.zeros(len(p))
    g.ter_prev, g.ter_cur, g.ter_motion, g.ter_order = i32(q, "prev"), i32(q, "cur"), i32(q, "motion"), i32(q, "order")
    g.ter_sigma, g.ter_huber = np.full(len(q), t["point_sigma"]), np.zeros(len(q))
    g.btw_a, g.btw_b, g.btw_order, g.btw_z = i32(b, "a"), i32(b, "b"), i32(b, "order"), f64(b, "z", 12)
    g.btw_sigmas = np.full((len(b), 6), t["pose_sigma"])
    g.pri_a, g.pri_order, g.pri_z = i32(pr, "a"), i32(pr, "order"), f64(pr, "z", 12)
    g.pri_sigmas = np.full((len(pr), 6), t["pose_sigma"])
    g.n_p2p, g.n_ternary, g.n_between, g.n_prior = len(p), len(q), len(b), len(pr)
    g.n_factors = g.n_p2p + g.n_ternary + g.n_between + g.n_prior
    g.pose_frames = lambda: np.array([d.key_decode(int(k))[2] for k in g.pose_keys], dtype=np.int32)
    return g


def numpy_hessian_pattern(g):
    """Block pattern of J^T J in natural order (poses, then points), lower triangle, CSR: straight from the key tuples."""
    import numpy as np
    n = g.n_pose + g.n_point
    pairs = [(i, i) for i in range(n)]
    pt = lambda j: g.n_pose + int(j)   # noqa: E731
    pairs += [(int(a), pt(b)) for a, b in zip(g.p2p_pose, g.p2p_point)]
    for a, b, h in zip(g.ter_prev, g.ter_cur, g.ter_motion):
        pairs += [(pt(a), pt(b)), (pt(a), int(h)), (pt(b), int(h))]
    pairs += [(int(a), int(b)) for a, b in zip(g.btw_a, g.btw_b)]
    rc = np.array(sorted({(max(a, b), min(a, b)) for a, b in pairs}), dtype=np.int64)
    row_ptr = np.zeros(n + 1, dtype=np.int64)
    np.add.at(row_ptr, rc[:, 0] + 1, 1)
    return np.cumsum(row_ptr), rc[:, 1].astype(np.int32)
