"""Post-solve extraction (dynosam_b200/postsolve.py) against the rules of the reference's propogateObjectPoses
(src/common/DynamicObjects.cc:48-201) and the CSV layout of its logger (src/logger/Logger.cc:151-172)."""
import csv

import numpy as np

import dynosam_b200 as d
from dynosam_b200 import postsolve as ps
import np_ref as ref


def _motion(seed):
    rng = np.random.default_rng(seed)
    return ref.se3_exp(np.concatenate([rng.normal(0, 0.05, 3), rng.normal(0, 0.4, 3)]))


def test_propagation_rules():
    H = {k: _motion(k) for k in range(1, 12)}
    motions = {3: {1: H[3]}, 4: {1: H[4]}, 6: {1: H[6]}, 10: {1: H[10]}, 11: {1: H[11], 2: H[2]}}
    cent = {(1, f): np.array([f, 0.5 * f, 1.0]) for f in range(0, 12)}
    cent.update({(2, f): np.array([-f, 2.0, 0.0]) for f in range(0, 12)})
    poses, how = ps.propagate_object_poses(motions, cent)
    p1 = poses[1]
    # new object at frame 3: identity rotation at the centroid of frame 2, then L_k = H_k L_{k-1}
    assert how[(1, 2)] == ps.INIT_CENTROID and np.allclose(p1[2][:3, :3], np.eye(3)) and np.allclose(p1[2][:3, 3], cent[(1, 2)])
    assert how[(1, 3)] == ps.PROPAGATE and np.allclose(p1[3], H[3] @ p1[2]) and np.allclose(p1[4], H[4] @ p1[3])
    # frame 6: no pose at 5, last pose at 4, gap 2 < 3: frames 4..6 interpolated towards (rotation of the last pose, centroid at 6)
    assert how[(1, 5)] == how[(1, 6)] == ps.INTERPOLATE
    assert np.allclose(p1[6][:3, :3], p1[4][:3, :3]) and np.allclose(p1[6][:3, 3], cent[(1, 6)])
    assert np.allclose(p1[5][:3, 3], 0.5 * (p1[4][:3, 3] + p1[6][:3, 3]), atol=1e-9)   # same rotation at both ends: the translation is linear
    # frame 10: last pose at 6, gap 4 >= 3: restart from the centroid at 9
    assert how[(1, 9)] == ps.INIT_CENTROID and np.allclose(p1[9][:3, 3], cent[(1, 9)]) and np.allclose(p1[10], H[10] @ p1[9])
    assert np.allclose(p1[11], H[11] @ p1[10])
    assert how[(2, 10)] == ps.INIT_CENTROID and np.allclose(poses[2][11], H[2] @ poses[2][10])
    assert 7 not in p1 and 8 not in p1


def test_extraction_on_simulator_scene(tmp_path):
    scene = d.Scene.simulator(20)
    g = scene.full_batch_graph()
    cents = ps.object_centroids(g, g.point_xyz)
    obj = ps.point_objects(g)
    chars = np.array([d.key_decode(int(k))[0] for k in g.point_keys])
    assert np.all((obj >= 0) == (chars == "m"))
    # a centroid is the mean of the object's points at that frame
    (o, f), c = next(iter(cents.items()))
    frames = np.array([d.key_decode(int(k))[2] for k in g.point_keys])
    assert np.allclose(c, g.point_xyz[(obj == o) & (frames == f)].mean(axis=0))
    # with the ground-truth motions in place of the values, chained poses follow the ground-truth object motion exactly
    gobj, gframe, gmot = scene.gt_motions()
    pose_Rt = g.pose_Rt.copy()
    index = {(int(o), int(f)): i for i, (o, f) in enumerate(zip(gobj, gframe))}
    for i, k in enumerate(g.pose_keys):
        c, o, f, _ = d.key_decode(int(k))
        if c == "H":
            pose_Rt[i] = gmot[index[(o, f)]]
    motions = ps.object_motions(g, pose_Rt)
    poses, how = ps.propagate_object_poses(motions, cents)
    for o, per in poses.items():
        fs = sorted(per)
        for f in fs[1:]:
            if how[(o, f)] == ps.PROPAGATE:
                assert np.allclose(per[f], ps._T(gmot[index[(o, f)]]) @ per[f - 1])
    paths = ps.write_logs(str(tmp_path), "rgbd_motion_world_backend", g, g.pose_Rt, g.point_xyz, scene.gt_camera(), scene.gt_motions())
    with open(paths["camera_pose_log.csv"]) as fh:
        rows = list(csv.reader(fh))
    assert rows[0] == ["frame_id", "tx", "ty", "tz", "qx", "qy", "qz", "qw", "gt_tx", "gt_ty", "gt_tz", "gt_qx", "gt_qy", "gt_qz", "gt_qw"]   # Logger.cc:151-153
    assert len(rows) == 1 + 20 and abs(sum(float(v) ** 2 for v in rows[3][4:8]) - 1.0) < 1e-12
    with open(paths["object_motion_log.csv"]) as fh:
        rows = list(csv.reader(fh))
    assert rows[0][:9] == ["frame_id", "object_id", "tx", "ty", "tz", "qx", "qy", "qz", "qw"] and len(rows) == 1 + sum(len(v) for v in motions.values())
    with open(paths["map_points_log.csv"]) as fh:
        assert next(csv.reader(fh)) == ["frame_id", "object_id", "tracklet_id", "x_world", "y_world", "z_world"]
