"""The reference's recorded front-end output, `rgbd_frontend_output.bson`.

What the file is. RGBDInstanceFrontendModule keeps every RGBDInstanceOutputPacket it produced in a
std::map<FrameId, RGBDInstanceOutputPacket::Ptr> and, at shutdown, JsonConverter::WriteOutJson wraps it as {"data": map}
and writes json::to_bson of that (reference src/frontend/RGBDInstanceFrontendModule.cc:80, 176;
include/dynosam/logger/Logger.hpp:193-231; file name include/dynosam/frontend/Frontend-Definitions.hpp:41).
FrontendOfflinePipeline replays the file into the back-end (include/dynosam/frontend/FrontendPipeline.hpp:54-87). The JSON
shape of a packet is fixed by the adl_serializers of include/dynosam/utils/JsonUtils.hpp and src/utils/JsonUtils.cc:68-118:

  data            [[frame_id, packet], ...]          std::map with a non-string key: nlohmann writes an array of pairs
  packet          frontend_type "RGB", frame_id, timestamp, T_world_camera, static_keypoints, dynamic_keypoints,
                  static_landmarks, dynamic_landmarks, estimated_motions, propogated_object_poses, camera_poses, ground_truth
  Pose3           {tx, ty, tz, qx, qy, qz, qw}                                            (JsonUtils.hpp:180-208)
  status          {value, frame_id, tracklet_id, object_id, reference_frame}              (JsonUtils.hpp:320-346)
  value           {measurement, covariance?}; a Point3 is an Eigen 3 x 1: [[x], [y], [z]]  (JsonUtils.hpp:150-176, 288-318)
  estimated_motions        [[object_id, {estimate: Pose3, reference_frame}], ...]          (gtsam::FastMap -> std::map)
  propogated_object_poses  [[object_id, [[frame_id, Pose3], ...]], ...]
  reference_frame "global" | "local" | "object"                                            (JsonUtils.hpp:51-55)

What the back-end reads from a packet is frame id, T_world_camera, the two landmark vectors (camera-frame measurements with
tracklet and object ids: the front-end fills them with LandmarkStatus::StaticInLocal / DynamicInLocal,
RGBDInstanceFrontendModule.cc:332-397) and the estimated motions (RGBDBackendModule::updateMap, RGBDBackendModule.cc:263-278);
that is what read_frontend_bson() turns into a Scene. Keypoints, ground truth, propagated poses and covariances are read past.
write_frontend_bson() writes the same shape from any Scene, with object keys in lexicographic order, as nlohmann's
std::map-backed objects come out.

The BSON codec below is the subset nlohmann::json emits and accepts (double 0x01, string 0x02, document 0x03, array 0x04,
binary 0x05 [read past], boolean 0x08, null 0x0A, int32 0x10, int64 0x12); integers are written as int32 when they fit, as
nlohmann does. No sample written by the reference exists in its repository (its own round-trip test, test/test_pipelines.cc:69-110,
generates one at run time), so this module is pinned by the BSON specification's known-answer documents, by the key names
above and by round trips through the library's scenes (tests/test_frontend_bson.py) -- not by a reference-written file.
"""
import struct

import numpy as np

REFERENCE_FRAMES = ("global", "local", "object")
PACKET_KEYS = ("T_world_camera", "camera_poses", "dynamic_keypoints", "dynamic_landmarks", "estimated_motions", "frame_id",
               "frontend_type", "ground_truth", "propogated_object_poses", "static_keypoints", "static_landmarks", "timestamp")


# ---------------------------------------------------------------------------------------------------------------- BSON
def _cstring(buf, at):
    end = buf.index(b"\x00", at)
    return buf[at:end].decode("utf-8"), end + 1


def _decode_document(buf, at, as_array):
    (size,) = struct.unpack_from("<i", buf, at)
    end = at + size
    if size < 5 or end > len(buf) or buf[end - 1] != 0:
        raise ValueError("BSON: bad document length at byte %d" % at)
    at += 4
    out = [] if as_array else {}
    while at < end - 1:
        tag = buf[at]
        name, at = _cstring(buf, at + 1)
        if tag == 0x01:
            (v,) = struct.unpack_from("<d", buf, at); at += 8
        elif tag == 0x02:
            (n,) = struct.unpack_from("<i", buf, at)
            v = buf[at + 4:at + 4 + n - 1].decode("utf-8"); at += 4 + n
        elif tag == 0x03:
            v, at = _decode_document(buf, at, False)
        elif tag == 0x04:
            v, at = _decode_document(buf, at, True)
        elif tag == 0x05:
            (n,) = struct.unpack_from("<i", buf, at)
            v = bytes(buf[at + 5:at + 5 + n]); at += 5 + n
        elif tag == 0x08:
            v = buf[at] != 0; at += 1
        elif tag == 0x0A:
            v = None
        elif tag == 0x10:
            (v,) = struct.unpack_from("<i", buf, at); at += 4
        elif tag == 0x12:
            (v,) = struct.unpack_from("<q", buf, at); at += 8
        else:
            raise ValueError("BSON: element type 0x%02x (key %r) is not one nlohmann::json writes" % (tag, name))
        if as_array:
            out.append(v)
        else:
            out[name] = v
    return out, end


def bson_decode(data):
    """bytes -> dict (json::from_bson)."""
    buf = bytes(data)
    doc, end = _decode_document(buf, 0, False)
    if end != len(buf):
        raise ValueError("BSON: %d trailing bytes" % (len(buf) - end))
    return doc


def _encode_element(name, v, out):
    key = name.encode("utf-8") + b"\x00"
    if v is None:
        out += b"\x0A" + key
    elif isinstance(v, (bool, np.bool_)):
        out += b"\x08" + key + (b"\x01" if v else b"\x00")
    elif isinstance(v, (int, np.integer)):
        v = int(v)
        if -(1 << 31) <= v < (1 << 31):
            out += b"\x10" + key + struct.pack("<i", v)
        elif -(1 << 63) <= v < (1 << 63):
            out += b"\x12" + key + struct.pack("<q", v)
        else:
            raise ValueError("BSON: integer %d does not fit int64 (nlohmann refuses it too)" % v)
    elif isinstance(v, (float, np.floating)):
        out += b"\x01" + key + struct.pack("<d", float(v))
    elif isinstance(v, str):
        b = v.encode("utf-8")
        out += b"\x02" + key + struct.pack("<i", len(b) + 1) + b + b"\x00"
    elif isinstance(v, dict):
        out += b"\x03" + key + _encode_document(v)
    elif isinstance(v, (list, tuple)):
        out += b"\x04" + key + _encode_document({str(i): x for i, x in enumerate(v)}, ordered=True)
    else:
        raise TypeError("BSON: cannot encode %r" % type(v))


def _encode_document(d, ordered=False):
    body = bytearray()
    for k in (d if ordered else sorted(d)):   # nlohmann's objects are std::map: keys in lexicographic order
        _encode_element(k, d[k], body)
    return struct.pack("<i", len(body) + 5) + bytes(body) + b"\x00"


def bson_encode(doc):
    """dict -> bytes (json::to_bson)."""
    if not isinstance(doc, dict):
        raise TypeError("BSON: the top level must be an object")
    return _encode_document(doc)


# ------------------------------------------------------------------------------------------------------------- schema
def pose_to_json(Rt):
    """12 doubles (row-major R, then t) -> the Pose3 object (Eigen quaternion of the rotation, qw >= 0 not enforced by the
    reference; written here with qw >= 0)."""
    R = np.asarray(Rt[:9], dtype=np.float64).reshape(3, 3)
    t = np.asarray(Rt[9:12], dtype=np.float64)
    tr = R[0, 0] + R[1, 1] + R[2, 2]
    if tr > 0:
        s = 2.0 * np.sqrt(tr + 1.0)
        q = np.array([(R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s, 0.25 * s])
    else:
        i = int(np.argmax([R[0, 0], R[1, 1], R[2, 2]]))
        j, k = (i + 1) % 3, (i + 2) % 3
        s = 2.0 * np.sqrt(1.0 + R[i, i] - R[j, j] - R[k, k])
        q = np.zeros(4)
        q[i] = 0.25 * s
        q[j] = (R[j, i] + R[i, j]) / s
        q[k] = (R[k, i] + R[i, k]) / s
        q[3] = (R[k, j] - R[j, k]) / s
    if q[3] < 0:
        q = -q
    return {"tx": float(t[0]), "ty": float(t[1]), "tz": float(t[2]), "qx": float(q[0]), "qy": float(q[1]), "qz": float(q[2]), "qw": float(q[3])}


def pose_from_json(j):
    """The Pose3 object -> 12 doubles; gtsam::Rot3(qw, qx, qy, qz) normalises the quaternion (Eigen toRotationMatrix of the
    normalised value)."""
    q = np.array([j["qx"], j["qy"], j["qz"], j["qw"]], dtype=np.float64)
    q = q / np.linalg.norm(q)
    x, y, z, w = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    return np.concatenate([R.reshape(-1), [float(j["tx"]), float(j["ty"]), float(j["tz"])]])


def _landmark_status(frame, tracklet, obj, z):
    return {"value": {"measurement": [[float(z[0])], [float(z[1])], [float(z[2])]]}, "frame_id": int(frame), "tracklet_id": int(tracklet),
            "object_id": int(obj), "reference_frame": "local"}


def _point_of(measurement):
    m = np.asarray(measurement, dtype=np.float64).reshape(-1)
    if m.shape != (3,):
        raise ValueError("landmark measurement is not a 3-vector")
    return m


def scene_to_json(scene):
    """{"data": [[frame_id, packet], ...]} of a Scene, in the reference's shape (keypoints, ground truth and propagated
    poses empty: the library's scenes do not carry them)."""
    pk = scene.packets()
    sf, st, sz = pk["static"]
    df, dt, do, dz = pk["dynamic"]
    mf, mo, mr = pk["motions"]
    s_lo, s_hi = np.searchsorted(sf, pk["frame_id"], "left"), np.searchsorted(sf, pk["frame_id"], "right")
    d_lo, d_hi = np.searchsorted(df, pk["frame_id"], "left"), np.searchsorted(df, pk["frame_id"], "right")
    m_lo, m_hi = np.searchsorted(mf, pk["frame_id"], "left"), np.searchsorted(mf, pk["frame_id"], "right")
    data, cams = [], []
    for i, f in enumerate(pk["frame_id"]):
        cam = pose_to_json(pk["T_world_camera"][i])
        cams.append(cam)
        packet = {
            "frontend_type": "RGB", "frame_id": int(f), "timestamp": float(f), "T_world_camera": cam,
            "static_keypoints": [], "dynamic_keypoints": [], "ground_truth": None,
            "static_landmarks": [_landmark_status(f, st[a], 0, sz[a]) for a in range(s_lo[i], s_hi[i])],
            "dynamic_landmarks": [_landmark_status(f, dt[a], do[a], dz[a]) for a in range(d_lo[i], d_hi[i])],
            "estimated_motions": [[int(mo[a]), {"estimate": pose_to_json(mr[a]), "reference_frame": "global"}] for a in range(m_lo[i], m_hi[i])],
            "propogated_object_poses": [], "camera_poses": list(cams),
        }
        data.append([int(f), packet])
    return {"data": data}


def json_to_packets(doc):
    """{"data": [[frame_id, packet], ...]} -> the flat arrays Scene.from_packets() takes. Packets are taken in ascending
    frame order (the map's order); a null packet (a null shared_ptr) is an error, as it is for the reference's reader."""
    rows = doc["data"]
    if isinstance(rows, dict):   # a map<string, packet> would arrive as an object; the reference's key is numeric, accept both
        rows = [[int(k), v] for k, v in rows.items()]
    rows = sorted(rows, key=lambda r: int(r[0]))
    fid, cam = [], []
    sf, st, sz, df, dt, do, dz, mf, mo, mr = ([] for _ in range(10))
    for f, p in rows:
        if p is None:
            raise ValueError("frame %d: null packet" % f)
        missing = [k for k in ("frame_id", "T_world_camera", "static_landmarks", "dynamic_landmarks", "estimated_motions") if k not in p]
        if missing:
            raise ValueError("frame %d: packet lacks %s" % (f, ", ".join(missing)))
        if int(p["frame_id"]) != int(f):
            raise ValueError("map key %d holds the packet of frame %d" % (f, p["frame_id"]))
        fid.append(int(f)); cam.append(pose_from_json(p["T_world_camera"]))
        for s in p["static_landmarks"] or []:
            if s.get("reference_frame", "local") != "local":
                raise ValueError("frame %d: static landmark in the %r frame (the back-end takes camera-frame measurements)" % (f, s["reference_frame"]))
            sf.append(int(s["frame_id"])); st.append(int(s["tracklet_id"])); sz.append(_point_of(s["value"]["measurement"]))
        for s in p["dynamic_landmarks"] or []:
            if s.get("reference_frame", "local") != "local":
                raise ValueError("frame %d: dynamic landmark in the %r frame" % (f, s["reference_frame"]))
            df.append(int(s["frame_id"])); dt.append(int(s["tracklet_id"])); do.append(int(s["object_id"])); dz.append(_point_of(s["value"]["measurement"]))
        for obj, est in p["estimated_motions"] or []:
            if est.get("reference_frame", "global") != "global":
                raise ValueError("frame %d: motion of object %d in the %r frame (the world-centric back-end takes ^w H)" % (f, obj, est["reference_frame"]))
            mf.append(int(f)); mo.append(int(obj)); mr.append(pose_from_json(est["estimate"]))
    arr = lambda a, dt_, shape: np.asarray(a, dtype=dt_).reshape(shape)
    return {"frame_id": arr(fid, np.int64, (-1,)), "T_world_camera": arr(cam, np.float64, (-1, 12)),
            "static": (arr(sf, np.int64, (-1,)), arr(st, np.int64, (-1,)), arr(sz, np.float64, (-1, 3))),
            "dynamic": (arr(df, np.int64, (-1,)), arr(dt, np.int64, (-1,)), arr(do, np.int32, (-1,)), arr(dz, np.float64, (-1, 3))),
            "motions": (arr(mf, np.int64, (-1,)), arr(mo, np.int32, (-1,)), arr(mr, np.float64, (-1, 12)))}


def write_frontend_bson(path, scene):
    with open(path, "wb") as f:
        f.write(bson_encode(scene_to_json(scene)))


def read_frontend_bson(path):
    """rgbd_frontend_output.bson -> Scene (JsonConverter::ReadInJson + the back-end's reading of each packet)."""
    from .api import Scene
    with open(path, "rb") as f:
        pk = json_to_packets(bson_decode(f.read()))
    return Scene.from_packets(pk["frame_id"], pk["T_world_camera"], pk["static"], pk["dynamic"], pk["motions"])
