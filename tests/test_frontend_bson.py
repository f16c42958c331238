"""The reference's recorded front-end output (rgbd_frontend_output.bson): the BSON codec against the
specification's known-answer documents, the packet shape against the key names of the reference's serialisers
(src/utils/JsonUtils.cc:68-118, include/dynosam/utils/JsonUtils.hpp), and scene -> file -> scene round trips that must give
the same factor graph. No reference-written sample exists in its repository (test/test_pipelines.cc:69-110 writes one at run
time), so these pins are the specification and the serialisers' source, not a reference file."""
import struct

import numpy as np
import pytest

import dynosam_b200 as d
from dynosam_b200 import frontend_bson as fb


def test_bson_known_answers():
    # bsonspec.org: {"hello": "world"} and {"BSON": ["awesome", 5.05, 1986]}
    hello = b"\x16\x00\x00\x00\x02hello\x00\x06\x00\x00\x00world\x00\x00"
    assert fb.bson_encode({"hello": "world"}) == hello
    assert fb.bson_decode(hello) == {"hello": "world"}
    awesome = (b"\x31\x00\x00\x00\x04BSON\x00\x26\x00\x00\x00\x02\x30\x00\x08\x00\x00\x00awesome\x00\x01\x31\x00\x33\x33\x33\x33\x33\x33\x14\x40"
               b"\x10\x32\x00\xc2\x07\x00\x00\x00\x00")
    assert fb.bson_encode({"BSON": ["awesome", 5.05, 1986]}) == awesome
    assert fb.bson_decode(awesome) == {"BSON": ["awesome", 5.05, 1986]}
    # nlohmann::json::to_bson documentation example: {"compact": true, "schema": 0}
    assert fb.bson_encode({"compact": True, "schema": 0}) == bytes([0x1b, 0, 0, 0, 0x08]) + b"compact\x00\x01" + b"\x10schema\x00" + bytes(4) + b"\x00"


def test_bson_types_and_key_order():
    doc = {"b": None, "a": {"z": 1 << 40, "y": -7, "x": [1.5, False, "s", [], {}]}, "c": 2 ** 31 - 1, "d": -(2 ** 31), "e": 2 ** 31}
    raw = fb.bson_encode(doc)
    assert fb.bson_decode(raw) == doc
    assert raw.index(b"\x03a\x00") < raw.index(b"\x0Ab\x00") < raw.index(b"\x10c\x00")      # std::map order
    assert b"\x10d\x00" + struct.pack("<i", -(2 ** 31)) in raw and b"\x12e\x00" + struct.pack("<q", 2 ** 31) in raw   # int32 when it fits
    with pytest.raises(ValueError):
        fb.bson_encode({"big": 1 << 63})   # nlohmann refuses uint64 beyond int64 as well
    with pytest.raises(ValueError):
        fb.bson_decode(raw[:-3])
    with pytest.raises(ValueError):
        fb.bson_decode(b"\x0c\x00\x00\x00\x07a\x00" + bytes(4) + b"\x00")   # ObjectId: not a type nlohmann writes


def test_pose_quaternion_round_trip():
    rng = np.random.default_rng(3)
    from scipy.spatial.transform import Rotation
    for R in list(Rotation.random(40, random_state=5).as_matrix()) + [np.eye(3), np.diag([1.0, -1.0, -1.0]), np.diag([-1.0, 1.0, -1.0]), np.diag([-1.0, -1.0, 1.0])]:
        Rt = np.concatenate([R.reshape(-1), rng.normal(size=3)])
        j = fb.pose_to_json(Rt)
        assert set(j) == {"tx", "ty", "tz", "qx", "qy", "qz", "qw"} and j["qw"] >= 0
        q = Rotation.from_matrix(R).as_quat()
        q = -q if q[3] < 0 else q
        if abs(q[3]) > 1e-9:
            assert np.allclose([j["qx"], j["qy"], j["qz"], j["qw"]], q, atol=1e-12)
        assert np.allclose(fb.pose_from_json(j), Rt, atol=1e-14)
    # gtsam::Rot3(w, x, y, z) normalises
    assert np.allclose(fb.pose_from_json({"tx": 0, "ty": 0, "tz": 0, "qx": 0, "qy": 0, "qz": 0, "qw": 2.0})[:9], np.eye(3).reshape(-1))


def test_packet_shape_follows_the_reference_serialisers():
    scene = d.Scene.simulator(n_frames=4)
    doc = fb.scene_to_json(scene)
    assert list(doc) == ["data"] and [r[0] for r in doc["data"]] == [0, 1, 2, 3]
    for f, p in doc["data"]:
        assert tuple(sorted(p)) == fb.PACKET_KEYS and p["frame_id"] == f and p["frontend_type"] == "RGB"
        assert len(p["camera_poses"]) == f + 1 and p["ground_truth"] is None
        for s in p["static_landmarks"] + p["dynamic_landmarks"]:
            assert set(s) == {"value", "frame_id", "tracklet_id", "object_id", "reference_frame"} and s["reference_frame"] == "local"
            m = s["value"]["measurement"]
            assert len(m) == 3 and all(len(row) == 1 for row in m)   # Eigen 3 x 1: rows of one column
        assert all(s["object_id"] == 0 for s in p["static_landmarks"]) and all(s["object_id"] >= 1 for s in p["dynamic_landmarks"])
        for obj, est in p["estimated_motions"]:
            assert set(est) == {"estimate", "reference_frame"} and est["reference_frame"] in fb.REFERENCE_FRAMES


def _same_graph(a, b, pose_tol):
    assert a.n_factors == b.n_factors
    for name in ("pose_keys", "point_keys", "p2p_pose", "p2p_point", "ter_prev", "ter_cur", "ter_motion", "btw_a", "btw_b", "pri_a"):
        assert np.array_equal(getattr(a, name), getattr(b, name)), name
    assert np.array_equal(a.p2p_z, b.p2p_z)                    # measurements travel as doubles: bit-exact
    for name in ("pose_Rt", "point_xyz", "btw_z", "pri_z"):   # poses pass through a quaternion
        assert np.allclose(getattr(a, name), getattr(b, name), atol=pose_tol), name


@pytest.mark.parametrize("make", [lambda: d.Scene.simulator(n_frames=12),
                                  lambda: d.Scene.synthetic(n_frames=30, n_objects=3, static_per_frame=40, dynamic_per_object=10, object_span=20, seed=11)])
def test_scene_file_round_trip_gives_the_same_graphs(tmp_path, make):
    scene = make()
    path = tmp_path / "rgbd_frontend_output.bson"
    fb.write_frontend_bson(path, scene)
    replay = fb.read_frontend_bson(path)
    pa, pb = scene.packets(), replay.packets()
    assert np.array_equal(pa["frame_id"], pb["frame_id"])
    for k in ("static", "dynamic"):
        for x, y in zip(pa[k], pb[k]):
            assert np.array_equal(x, y)
    assert np.array_equal(pa["motions"][0], pb["motions"][0]) and np.array_equal(pa["motions"][1], pb["motions"][1])
    assert np.allclose(pa["motions"][2], pb["motions"][2], atol=1e-13) and np.allclose(pa["T_world_camera"], pb["T_world_camera"], atol=1e-13)
    _same_graph(scene.full_batch_graph(), replay.full_batch_graph(), 1e-12)
    _same_graph(scene.window_graph(2, 9, True), replay.window_graph(2, 9, True), 1e-12)
    # a second write of the replayed scene is byte-identical to a write of what was read (the codec is a fixed point)
    again = tmp_path / "again.bson"
    fb.write_frontend_bson(again, replay)
    assert fb.bson_decode(again.read_bytes()) == fb.scene_to_json(replay)


def test_reader_accepts_what_the_reference_adds_and_rejects_what_the_backend_cannot_use():
    pose = {"tx": 1.0, "ty": 2.0, "tz": 3.0, "qx": 0.0, "qy": 0.0, "qz": 0.0, "qw": 1.0}
    status = lambda f, t, o, z: {"value": {"measurement": [[z[0]], [z[1]], [z[2]]], "covariance": [[1, 0, 0], [0, 1, 0], [0, 0, 1]]},
                                 "frame_id": f, "tracklet_id": t, "object_id": o, "reference_frame": "local"}
    def packet(f):
        return {"frontend_type": "RGB", "frame_id": f, "timestamp": 0.1 * f, "T_world_camera": pose, "static_keypoints": [{"ignored": 1}], "dynamic_keypoints": [],
                "static_landmarks": [status(f, 5, 0, (1.0, 0.0, 4.0)), status(f, 6, 0, (0.5, 0.5, 3.0))], "dynamic_landmarks": [status(f, 9, 2, (0.0, 1.0, 6.0))],
                "estimated_motions": [[2, {"estimate": pose, "reference_frame": "global"}]] if f else [], "propogated_object_poses": [[2, [[f, pose]]]],
                "camera_poses": [pose], "ground_truth": {"timestamp": 0.0, "frame_id": f, "X_world": pose, "objects": []}}
    doc = {"data": [[7, packet(7)], [5, packet(5)], [6, packet(6)]]}   # any order in the file: the map sorts
    pk = fb.json_to_packets(fb.bson_decode(fb.bson_encode(doc)))
    assert list(pk["frame_id"]) == [5, 6, 7] and len(pk["static"][0]) == 6 and list(pk["dynamic"][2]) == [2, 2, 2]
    assert list(pk["motions"][0]) == [5, 6, 7] and list(pk["motions"][1]) == [2, 2, 2]
    scene = d.Scene.from_packets(pk["frame_id"], pk["T_world_camera"], pk["static"], pk["dynamic"], pk["motions"])
    assert scene.info() == {"n_frames": 3, "n_static_obs": 6, "n_dynamic_obs": 3}
    with pytest.raises(RuntimeError):
        scene.gt_camera()   # recorded packets carry no ground truth the library could hand back
    bad = {"data": [[0, dict(packet(0), static_landmarks=[dict(status(0, 1, 0, (0, 0, 1)), reference_frame="global")])]]}
    with pytest.raises(ValueError):
        fb.json_to_packets(bad)
    with pytest.raises(ValueError):
        fb.json_to_packets({"data": [[0, None]]})
    with pytest.raises(ValueError):
        fb.json_to_packets({"data": [[3, packet(4)]]})
    with pytest.raises(RuntimeError):   # an observation of a frame without a packet
        identity = np.concatenate([np.eye(3).reshape(-1), np.zeros(3)]).reshape(1, 12)
        d.Scene.from_packets([0], identity, ([1], [0], [[0, 0, 1]]), ([], [], [], np.zeros((0, 3))), ([], [], np.zeros((0, 12))))


@pytest.mark.gpu
def test_replayed_file_solves_like_the_scene_it_came_from(tmp_path):
    scene = d.Scene.synthetic(n_frames=40, n_objects=3, static_per_frame=60, dynamic_per_object=12, object_span=25, seed=4)
    path = tmp_path / "rgbd_frontend_output.bson"
    fb.write_frontend_bson(path, scene)
    replay = fb.read_frontend_bson(path)
    out = []
    for s in (scene, replay):
        lm = d.GpuLevenbergMarquardt(s.full_batch_graph(), d.default_lm_params())
        lm.optimize()
        out.append((lm.error(), lm.iterations()))
    assert out[0][1] == out[1][1]
    assert abs(out[0][0] - out[1][0]) <= 1e-9 * abs(out[0][0])   # initial poses passed through a quaternion: round-off only
