"""Regenerates three_frame.json:  python tests/golden/make_three_frame.py

The one fully specified graph the reference's own tests hold for this path: the hand-built world-motion graph of
reference dynosam/test/test_rgbd_backend.cc:272-664 (TEST testCliques): a static landmark seen from three (then four)
camera poses, three dynamic tracklets of one object over frames 0..2 (then 0..3), odometry, a prior on the first pose.
The reference test asserts no output (it draws the Bayes tree), so this is a golden INPUT: values, factors, keys and
insertion order restated here in plain numpy (Rodrigues / Expmap by scipy), with the key encoding written out again
(Symbol = chr << 56 | index, LabeledSymbol = chr << 56 | label << 48 | index, Cantor pairing) so that the library's
key functions are checked against it. Noise: Isotropic sigma 10 on every point factor (:300), Isotropic sigma 0.1 on the
pose factors (:445); no robust kernel.

"graph" is the first graph (frames 0..2, :302-495); "full_graph" the one after the update (frames 0..3, :537-600).
"""
import json
import os

import numpy as np
from scipy.spatial.transform import Rotation

HERE = os.path.dirname(os.path.abspath(__file__))


def symbol(c, j):
    return (ord(c) << 56) | j


def labeled(c, label, j):
    return (ord(c) << 56) | (label << 48) | j


def cantor(a, b):
    return (a + b) * (a + b + 1) // 2 + b


def X(k): return symbol("X", k)                                # CameraPoseSymbol, BackendDefinitions.hpp:56-60
def l(t): return symbol("l", t)                                # StaticLandmarkSymbol
def m(frame, t): return symbol("m", cantor(t, frame))          # DynamicLandmarkSymbol(frame, tracklet), DynamicPointSymbol.cc:31-44
def H(obj, k): return labeled("H", obj + ord("0"), k)          # ObjectMotionSymbol(object, frame), BackendDefinitions.hpp:76-83


def pose(rotvec, t):
    T = np.eye(4)
    T[:3, :3] = Rotation.from_rotvec(rotvec).as_matrix()
    T[:3, 3] = t
    return T


def Rt(T):
    return np.concatenate([T[:3, :3].reshape(-1), T[:3, 3]]).tolist()


def apply(T, p):
    return T[:3, :3] @ np.asarray(p, float) + T[:3, 3]


def apply_inv(T, p):
    return T[:3, :3].T @ (np.asarray(p, float) - T[:3, 3])


def build(n_frames):
    H01, H12, H23 = pose([0.3, 0.2, 0.1], [0, 1, 1]), pose([0.5, 0.1, 0.1], [1, 1.5, 1]), pose([0.4, 0.15, 0.1], [0, 3, 1])
    H0 = [np.eye(4), H01, H01 @ H12, H01 @ H12 @ H23]          # H_0_k
    dyn = [np.array([2, 1, 3.0]), np.array([1, 1, 3.0]), np.array([3, 0.5, 2.0])]   # tracklets 1, 2, 3 at frame 0
    cam = [pose([0.2, 0.1, 0.12], [1, 3, 4]), pose([0.1, 0.2, 1.570796], [1, 2, 1]), pose([0.2, 0.3, 3.141593], [2, 2, 1]),
           pose([0.1, 0.4, 4.712389], [-1, 1, 0])]
    static_obs = [[1, 2, 3], [2, 2, 3], [3, 2, 3], [4, 2, 3]]
    factors = []   # insertion order
    values = {}

    def first_graph():
        for k in range(3):
            factors.append(("p2p", X(k), l(0), static_obs[k]))
        for t in (1, 2, 3):
            for k in (1, 2):
                factors.append(("ternary", m(k - 1, t), m(k, t), H(1, k)))
        for t in (1, 2, 3):
            for k in range(3):
                factors.append(("p2p", X(k), m(k, t), apply_inv(cam[k], apply(H0[k], dyn[t - 1])).tolist()))
        for k in (1, 2):
            factors.append(("between", X(k - 1), X(k), Rt(np.linalg.inv(cam[k - 1]) @ cam[k])))
        factors.append(("prior", X(0), Rt(cam[0])))
        values[l(0)] = [0.0, 1.0, 1.0]
        for k in range(3):
            for t in (1, 2, 3):
                values[m(k, t)] = apply(H0[k], dyn[t - 1]).tolist()
        for k in (1, 2):
            values[H(1, k)] = Rt(np.eye(4))
        for k in range(3):
            values[X(k)] = Rt(cam[k])

    def update():
        for t in (1, 2, 3):
            factors.append(("ternary", m(2, t), m(3, t), H(1, 3)))
        for t in (1, 2, 3):
            factors.append(("p2p", X(3), m(3, t), apply_inv(cam[3], apply(H0[3], dyn[t - 1])).tolist()))
        factors.append(("p2p", X(3), l(0), static_obs[3]))
        factors.append(("between", X(2), X(3), Rt(np.linalg.inv(cam[2]) @ cam[3])))
        for t in (1, 2, 3):
            values[m(3, t)] = apply(H0[3], dyn[t - 1]).tolist()
        values[H(1, 3)] = Rt(np.eye(4))
        values[X(3)] = Rt(cam[3])

    first_graph()
    if n_frames == 4:
        update()
    pose_keys = sorted(k for k, v in values.items() if len(v) == 12)
    point_keys = sorted(k for k, v in values.items() if len(v) == 3)
    pi = {k: i for i, k in enumerate(pose_keys)}
    qi = {k: i for i, k in enumerate(point_keys)}
    g = {"pose_keys": [str(k) for k in pose_keys], "pose_Rt": [values[k] for k in pose_keys],
         "point_keys": [str(k) for k in point_keys], "point_xyz": [values[k] for k in point_keys],
         "point_sigma": 10.0, "pose_sigma": 0.1,
         "p2p": [], "ternary": [], "between": [], "prior": []}
    for order, f in enumerate(factors):
        if f[0] == "p2p":
            g["p2p"].append({"order": order, "pose": pi[f[1]], "point": qi[f[2]], "z": list(map(float, f[3]))})
        elif f[0] == "ternary":
            g["ternary"].append({"order": order, "prev": qi[f[1]], "cur": qi[f[2]], "motion": pi[f[3]]})
        elif f[0] == "between":
            g["between"].append({"order": order, "a": pi[f[1]], "b": pi[f[2]], "z": f[3]})
        else:
            g["prior"].append({"order": order, "a": pi[f[1]], "z": f[2]})
    return g


if __name__ == "__main__":
    out = {"source": "reference dynosam/test/test_rgbd_backend.cc:272-664 (testCliques)", "graph": build(3), "full_graph": build(4),
           "known_keys": {"X0": str(X(0)), "l0": str(l(0)), "m_frame2_tracklet3": str(m(2, 3)), "H_object1_frame2": str(H(1, 2))}}
    with open(os.path.join(HERE, "three_frame.json"), "w") as f:
        json.dump(out, f, indent=1)
    # the same first graph as C arrays for tests/abi_smoke.c (a GTSAM-free C client of the C ABI)
    g = out["graph"]

    def arr(ctype, name, values):
        flat = np.asarray(values).reshape(-1).tolist()
        fmt = "%.17g" if ctype == "double" else "%d"
        body = ", ".join((fmt % v) if ctype != "uint64_t" else ("%dULL" % int(v)) for v in flat)
        return "static const %s %s[] = {%s};\n" % (ctype, name, body)

    with open(os.path.join(HERE, "three_frame.h"), "w") as f:
        f.write("/* generated by tests/golden/make_three_frame.py from reference dynosam/test/test_rgbd_backend.cc:272-495; do not edit */\n")
        f.write("#include <stdint.h>\n")
        f.write("enum { TF_N_POSE = %d, TF_N_POINT = %d, TF_N_P2P = %d, TF_N_TER = %d, TF_N_BTW = %d, TF_N_PRI = %d };\n"
                % (len(g["pose_keys"]), len(g["point_keys"]), len(g["p2p"]), len(g["ternary"]), len(g["between"]), len(g["prior"])))
        f.write("static const double TF_POINT_SIGMA = %.17g, TF_POSE_SIGMA = %.17g;\n" % (g["point_sigma"], g["pose_sigma"]))
        f.write(arr("uint64_t", "tf_pose_keys", [int(k) for k in g["pose_keys"]]))
        f.write(arr("double", "tf_pose_Rt", g["pose_Rt"]))
        f.write(arr("uint64_t", "tf_point_keys", [int(k) for k in g["point_keys"]]))
        f.write(arr("double", "tf_point_xyz", g["point_xyz"]))
        f.write(arr("int32_t", "tf_p2p_pose", [r["pose"] for r in g["p2p"]]) + arr("int32_t", "tf_p2p_point", [r["point"] for r in g["p2p"]]))
        f.write(arr("int32_t", "tf_p2p_order", [r["order"] for r in g["p2p"]]) + arr("double", "tf_p2p_z", [r["z"] for r in g["p2p"]]))
        f.write(arr("int32_t", "tf_ter_prev", [r["prev"] for r in g["ternary"]]) + arr("int32_t", "tf_ter_cur", [r["cur"] for r in g["ternary"]]))
        f.write(arr("int32_t", "tf_ter_motion", [r["motion"] for r in g["ternary"]]) + arr("int32_t", "tf_ter_order", [r["order"] for r in g["ternary"]]))
        f.write(arr("int32_t", "tf_btw_a", [r["a"] for r in g["between"]]) + arr("int32_t", "tf_btw_b", [r["b"] for r in g["between"]]))
        f.write(arr("int32_t", "tf_btw_order", [r["order"] for r in g["between"]]) + arr("double", "tf_btw_z", [r["z"] for r in g["between"]]))
        f.write(arr("int32_t", "tf_pri_a", [r["a"] for r in g["prior"]]) + arr("int32_t", "tf_pri_order", [r["order"] for r in g["prior"]]))
        f.write(arr("double", "tf_pri_z", [r["z"] for r in g["prior"]]))
    print("wrote three_frame.json:", {k: len(out["graph"][k]) for k in ("pose_keys", "point_keys", "p2p", "ternary", "between", "prior")})
