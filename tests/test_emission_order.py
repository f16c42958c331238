"""Graph emission pinned beyond counts: factor insertion order, per-factor key tuples and the initial
values, checked from the flattened graph's `order` arrays with plain numpy.

Reference order (Formulation-impl.hpp:103,157,296,581; RGBDBackendModule.cc:165-200): per frame k -- odometry
BetweenFactor(X_{k-1}, X_k), then the static PoseToPoint factors in tracklet-id order, then per object (id order) per
tracklet [PoseToPoint at k-1 when the tracklet starts,] PoseToPoint at k, LandmarkMotionTernary(m^{k-1}, m^k, H^k), then
the motion-smoothing BetweenFactor(H^{k-1}, H^k). Quirks: frame 0 only holds the prior; a starting dynamic point at k is
initialised with the camera pose of k-1 (Formulation-impl.hpp:460-478); H starts at identity."""
import numpy as np
import pytest

import dynosam_b200 as d
import scenes


def _decode(keys):
    out = [d.key_decode(int(k)) for k in keys]
    return np.array([o[0] for o in out]), np.array([o[1] for o in out]), np.array([o[2] for o in out]), np.array([o[3] for o in out])


def _apply(Rt, z):
    return Rt[:9].reshape(3, 3) @ z + Rt[9:]


@pytest.mark.parametrize("name", ["simulator_cfg1", "small_noisy"])
def test_insertion_order_and_initial_values(name):
    g = (scenes.simulator_cfg1() if name == "simulator_cfg1" else scenes.small_noisy()).full_batch_graph()
    pchr, pobj, pframe, _ = _decode(g.pose_keys)
    qchr, _, qframe, qtrack = _decode(g.point_keys)
    n = g.n_factors
    kind, frame, obj, track = np.zeros(n, int), np.zeros(n, int), np.full(n, -1), np.full(n, -1)
    # kind ranks inside a frame: 0 prior / odometry, 1 static point, 2 dynamic point factors, 3 smoothing
    o = g.p2p_order
    dyn = qchr[g.p2p_point] == "m"
    kind[o] = np.where(dyn, 2, 1); frame[o] = np.where(dyn, qframe[g.p2p_point], pframe[g.p2p_pose]); track[o] = qtrack[g.p2p_point]
    assert np.all(pchr[g.p2p_pose] == "X")
    assert np.all(qframe[g.p2p_point][dyn] == pframe[g.p2p_pose][dyn])            # a dynamic point is observed from its own frame's camera
    o = g.ter_order
    assert np.all(pchr[g.ter_motion] == "H") and np.all(qchr[g.ter_prev] == "m") and np.all(qchr[g.ter_cur] == "m")
    assert np.all(qframe[g.ter_cur] == pframe[g.ter_motion]) and np.all(qframe[g.ter_prev] + 1 == qframe[g.ter_cur])   # (m^{k-1}, m^k, H^k)
    assert np.all(qtrack[g.ter_prev] == qtrack[g.ter_cur])
    kind[o] = 2; frame[o] = pframe[g.ter_motion]; obj[o] = pobj[g.ter_motion]; track[o] = qtrack[g.ter_cur]
    o = g.btw_order
    odo = pchr[g.btw_a] == "X"
    assert np.all(pchr[g.btw_a] == pchr[g.btw_b]) and np.all(pframe[g.btw_a] + 1 == pframe[g.btw_b])
    assert np.all(pobj[g.btw_a][~odo] == pobj[g.btw_b][~odo])
    kind[o] = np.where(odo, 0, 3); frame[o] = pframe[g.btw_b]; obj[o] = pobj[g.btw_b]
    kind[g.pri_order] = 0; frame[g.pri_order] = pframe[g.pri_a]
    assert sorted(np.concatenate([g.p2p_order, g.ter_order, g.btw_order, g.pri_order]).tolist()) == list(range(n))
    assert g.n_prior == 1 and g.pri_order[0] == 0 and pframe[g.pri_a[0]] == pframe[pchr == "X"].min()
    # a starting tracklet's first PoseToPoint factor is the observation at k-1, emitted while frame k is processed
    first_of_track = {}
    for i in np.argsort(g.p2p_order):
        if dyn[i]:
            first_of_track.setdefault(int(qtrack[g.p2p_point[i]]), i)
    emit_frame = frame.copy()
    for i in first_of_track.values():
        emit_frame[g.p2p_order[i]] += 1
    assert np.all(np.diff(emit_frame) >= 0)                                        # frames in time order
    for k in np.unique(emit_frame):
        sel = np.flatnonzero(emit_frame == k)
        assert np.all(np.diff(kind[sel]) >= 0), k                                  # odometry, static, dynamic, smoothing
        st = sel[kind[sel] == 1]
        assert np.all(np.diff(track[st]) > 0)                                      # static factors in tracklet-id order
    # dynamic factors of a tracklet: p2p (k-1 when starting), p2p k, ternary -- consecutive slots
    tcur_order = dict(zip(g.ter_cur.tolist(), g.ter_order.tolist()))
    p2p_order_of_point = dict(zip(g.p2p_point[dyn].tolist(), g.p2p_order[dyn].tolist()))
    for cur, prev in zip(g.ter_cur.tolist(), g.ter_prev.tolist()):
        assert tcur_order[cur] == p2p_order_of_point[cur] + 1
        if prev not in tcur_order:                                                 # the tracklet starts here
            assert p2p_order_of_point[prev] == p2p_order_of_point[cur] - 1
    # initial values
    assert np.allclose(g.pose_Rt[pchr == "H"], np.tile(np.r_[np.eye(3).reshape(-1), np.zeros(3)], ((pchr == "H").sum(), 1)))
    cam_of_frame = {int(f): i for i, f in enumerate(pframe) if pchr[i] == "X"}
    seen_static = set()
    for i in np.argsort(g.p2p_order):
        pt, z = int(g.p2p_point[i]), g.p2p_z[i]
        if not dyn[i]:
            if pt not in seen_static:                                              # first factor initialises the landmark from its own pose
                assert np.allclose(g.point_xyz[pt], _apply(g.pose_Rt[g.p2p_pose[i]], z), rtol=0, atol=1e-12)
                seen_static.add(pt)
            continue
        k = int(qframe[pt])
        starting_pair = pt in set(g.ter_prev.tolist()) and pt not in tcur_order   # the k-1 point of a starting pair
        second_of_start = pt in tcur_order and int(g.ter_prev[np.flatnonzero(g.ter_cur == pt)[0]]) not in tcur_order
        pose_frame = k - 1 if second_of_start else k                               # the quirk: X_{k-1} initialises m^k when the tracklet starts
        assert np.allclose(g.point_xyz[pt], _apply(g.pose_Rt[cam_of_frame[pose_frame]], z), rtol=0, atol=1e-12), (k, starting_pair)
