"""The DynoSAM text graph file (interchange format on the input side of the hot path).

Layout as written by the reference's NonlinearFactorGraphManager::writeDynosamGraphFile
(src/backend/FactorGraphTools.cc:256-381, helpers include/dynosam/backend/FactorGraphTools.hpp:414-438 and
include/dynosam/utils/Numerical.hpp:276-294): one line per factor, followed by one line for every variable
that factor is the first to mention:

    SE3_PRIOR_FACTOR     key            x y z qx qy qz qw   <21 upper-triangular information entries>
    SE3_BETWEEN_FACTOR   key1 key2      x y z qx qy qz qw   <21>
    SE3_MOTION_FACTOR    prev cur H     0 0 0               <6>        (LandmarkMotionTernaryFactor)
    POSE_TO_POINT_FACTOR pose point     zx zy zz            <6>
    SE3_POSE_VALUE / SE3_MOTION_VALUE        key   x y z qx qy qz qw
    POINT3_STATIC_VALUE / POINT3_DYNAMIC_VALUE  key   x y z

Keys are the 64-bit gtsam keys in decimal. The information matrix is that of the Gaussian part of the noise model;
a robust (Huber) wrapper is not recorded in the file, so the reader takes the Huber constant as an argument.
Nothing here touches the GPU: read_graph_file returns a HostGraph whose arrays go through
GpuLevenbergMarquardt.load_host_arrays (the C ABI's add_*_v calls)."""
import numpy as np

from .api import key_decode

FACTOR_TAGS = {"SE3_PRIOR_FACTOR": 1, "SE3_BETWEEN_FACTOR": 2, "SE3_MOTION_FACTOR": 3, "POSE_TO_POINT_FACTOR": 2}
POSE_VALUE_TAGS = ("SE3_POSE_VALUE", "SE3_MOTION_VALUE")
POINT_VALUE_TAGS = ("POINT3_STATIC_VALUE", "POINT3_DYNAMIC_VALUE")


def quat_to_R(qx, qy, qz, qw):
    n = np.sqrt(qx * qx + qy * qy + qz * qz + qw * qw)
    x, y, z, w = qx / n, qy / n, qz / n, qw / n
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def R_to_quat(R):
    """(qx, qy, qz, qw) with qw >= 0 (Eigen's Quaternion(Matrix3) branch structure)."""
    t = np.trace(R)
    if t > 0:
        s = 2.0 * np.sqrt(t + 1.0)
        q = ((R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s, 0.25 * s)
    else:
        i = int(np.argmax(np.diag(R)))
        j, k = (i + 1) % 3, (i + 2) % 3
        s = 2.0 * np.sqrt(R[i, i] - R[j, j] - R[k, k] + 1.0)
        v = [0.0, 0.0, 0.0]
        v[i] = 0.25 * s
        v[j] = (R[j, i] + R[i, j]) / s
        v[k] = (R[k, i] + R[i, k]) / s
        q = (v[0], v[1], v[2], (R[k, j] - R[j, k]) / s)
    q = np.array(q)
    return q if q[3] >= 0 else -q


def _pose_fields(Rt, fmt):
    R = np.asarray(Rt[:9]).reshape(3, 3)
    q = R_to_quat(R)
    return " ".join(fmt % v for v in (Rt[9], Rt[10], Rt[11], q[0], q[1], q[2], q[3]))


def _info_fields(sigmas, fmt):
    """upper triangle (row-major, i <= j) of diag(1 / sigma^2), each entry preceded by a blank as in the reference"""
    n = len(sigmas)
    out = []
    for i in range(n):
        for j in range(i, n):
            out.append(fmt % (1.0 / (sigmas[i] * sigmas[i]) if i == j else 0.0))
    return "".join(" " + v for v in out)


class HostGraph:
    """Flattened graph held in caller-owned numpy arrays, attribute-compatible with dynosam_b200.Graph."""

    def pose_frames(self):
        return np.array([key_decode(int(k))[2] for k in self.pose_keys], dtype=np.int32)


def write_graph_file(path, g, precision=17):
    """Writes g (Graph or HostGraph) in the reference's layout, factors in the caller's insertion order
    (g.*_order when present). precision=6 reproduces the reference's default ostream formatting."""
    fmt = "%%.%dg" % precision
    lines = []
    if all(hasattr(g, a) for a in ("p2p_order", "ter_order", "btw_order", "pri_order")) and g.n_factors and min(
            [int(o.min()) for o in (g.p2p_order, g.ter_order, g.btw_order, g.pri_order) if len(o)]) >= 0:
        slots = [(int(o), "p2p", i) for i, o in enumerate(g.p2p_order)] + [(int(o), "ter", i) for i, o in enumerate(g.ter_order)] + \
                [(int(o), "btw", i) for i, o in enumerate(g.btw_order)] + [(int(o), "pri", i) for i, o in enumerate(g.pri_order)]
        slots.sort()
    else:
        slots = [(0, "pri", i) for i in range(g.n_prior)] + [(0, "btw", i) for i in range(g.n_between)] + \
                [(0, "p2p", i) for i in range(g.n_p2p)] + [(0, "ter", i) for i in range(g.n_ternary)]
    seen = set()

    def value_line(kind, idx):
        key = int(g.pose_keys[idx]) if kind == "pose" else int(g.point_keys[idx])
        if key in seen:
            return
        seen.add(key)
        c = key_decode(key)[0]
        if kind == "pose":
            tag = "SE3_POSE_VALUE" if c == "X" else "SE3_MOTION_VALUE"
            lines.append("%s %d %s" % (tag, key, _pose_fields(g.pose_Rt[idx], fmt)))
        else:
            tag = "POINT3_STATIC_VALUE" if c == "l" else "POINT3_DYNAMIC_VALUE"
            lines.append("%s %d %s" % (tag, key, " ".join(fmt % v for v in g.point_xyz[idx])))

    for _, kind, i in slots:
        if kind == "pri":
            a = int(g.pri_a[i])
            lines.append("SE3_PRIOR_FACTOR %d %s%s" % (int(g.pose_keys[a]), _pose_fields(g.pri_z[i], fmt), _info_fields(g.pri_sigmas[i], fmt)))
            value_line("pose", a)
        elif kind == "btw":
            a, b = int(g.btw_a[i]), int(g.btw_b[i])
            lines.append("SE3_BETWEEN_FACTOR %d %d %s%s" % (int(g.pose_keys[a]), int(g.pose_keys[b]), _pose_fields(g.btw_z[i], fmt), _info_fields(g.btw_sigmas[i], fmt)))
            value_line("pose", a); value_line("pose", b)
        elif kind == "ter":
            a, b, h = int(g.ter_prev[i]), int(g.ter_cur[i]), int(g.ter_motion[i])
            lines.append("SE3_MOTION_FACTOR %d %d %d %s%s" % (int(g.point_keys[a]), int(g.point_keys[b]), int(g.pose_keys[h]), "0 0 0", _info_fields([g.ter_sigma[i]] * 3, fmt)))
            value_line("point", a); value_line("point", b); value_line("pose", h)
        else:
            a, b = int(g.p2p_pose[i]), int(g.p2p_point[i])
            lines.append("POSE_TO_POINT_FACTOR %d %d %s%s" % (int(g.pose_keys[a]), int(g.point_keys[b]), " ".join(fmt % v for v in g.p2p_z[i]), _info_fields([g.p2p_sigma[i]] * 3, fmt)))
            value_line("pose", a); value_line("point", b)
    with open(path, "w") as f:
        f.write("\n".join(lines) + ("\n" if lines else ""))


def _sigmas_from_info(vals, n):
    """sigmas of a diagonal information matrix stored as its upper triangle; off-diagonal entries must vanish"""
    sig, at = [], 0
    for i in range(n):
        for j in range(i, n):
            v = vals[at]; at += 1
            if i == j:
                if not v > 0:
                    raise ValueError("information matrix with a non-positive diagonal entry")
                sig.append(1.0 / np.sqrt(v))
            elif v != 0.0:
                raise ValueError("only diagonal noise models are part of the DynoSAM back-end (off-diagonal information %g)" % v)
    return sig


def read_graph_file(path, huber_k=0.0):
    """Parses a DynoSAM graph file into a HostGraph. huber_k > 0 wraps every point factor (PoseToPoint, motion ternary)
    in the Huber kernel the back-end applies when --use_robust_kernels is on (RGBDBackendModule.cc:97-113)."""
    pose_val, point_val = {}, {}
    pri, btw, ter, p2p = [], [], [], []
    slot = 0
    with open(path) as f:
        for ln, line in enumerate(f, 1):
            tok = line.split()
            if not tok:
                continue
            tag = tok[0]
            try:
                if tag in FACTOR_TAGS:
                    nk = FACTOR_TAGS[tag]
                    keys = [int(t) for t in tok[1:1 + nk]]
                    vals = [float(t) for t in tok[1 + nk:]]
                    if tag in ("SE3_PRIOR_FACTOR", "SE3_BETWEEN_FACTOR"):
                        if len(vals) != 7 + 21:
                            raise ValueError("expected 7 pose and 21 information entries, got %d" % len(vals))
                        Rt = np.concatenate([quat_to_R(*vals[3:7]).reshape(-1), vals[0:3]])
                        (pri if nk == 1 else btw).append((slot, keys, Rt, _sigmas_from_info(vals[7:], 6)))
                    else:
                        if len(vals) != 3 + 6:
                            raise ValueError("expected 3 measurement and 6 information entries, got %d" % len(vals))
                        sig = _sigmas_from_info(vals[3:], 3)
                        if not (sig[0] == sig[1] == sig[2]):
                            raise ValueError("point factors carry isotropic noise in the DynoSAM back-end")
                        (ter if nk == 3 else p2p).append((slot, keys, vals[0:3], sig[0]))
                    slot += 1
                elif tag in POSE_VALUE_TAGS:
                    v = [float(t) for t in tok[2:9]]
                    pose_val[int(tok[1])] = np.concatenate([quat_to_R(*v[3:7]).reshape(-1), v[0:3]])
                elif tag in POINT_VALUE_TAGS:
                    point_val[int(tok[1])] = [float(t) for t in tok[2:5]]
                else:
                    raise ValueError("unknown tag %r" % tag)
            except (ValueError, IndexError) as e:
                raise ValueError("%s:%d: %s" % (path, ln, e)) from None
    g = HostGraph()
    g.pose_keys = np.array(sorted(pose_val), dtype=np.uint64)          # gtsam::Values order: ascending key
    g.point_keys = np.array(sorted(point_val), dtype=np.uint64)
    g.pose_Rt = np.array([pose_val[int(k)] for k in g.pose_keys], dtype=float).reshape(-1, 12)
    g.point_xyz = np.array([point_val[int(k)] for k in g.point_keys], dtype=float).reshape(-1, 3)
    pidx = {int(k): i for i, k in enumerate(g.pose_keys)}
    qidx = {int(k): i for i, k in enumerate(g.point_keys)}

    def look(table, key, what):
        if key not in table:
            raise KeyError("factor on %s key %d that has no value line (gtsam::ValuesKeyDoesNotExist)" % (what, key))
        return table[key]

    g.n_pose, g.n_point = len(g.pose_keys), len(g.point_keys)
    g.n_p2p, g.n_ternary, g.n_between, g.n_prior = len(p2p), len(ter), len(btw), len(pri)
    g.n_factors = slot
    g.p2p_pose = np.array([look(pidx, k[0], "pose") for _, k, _, _ in p2p], dtype=np.int32)
    g.p2p_point = np.array([look(qidx, k[1], "point") for _, k, _, _ in p2p], dtype=np.int32)
    g.p2p_z = np.array([z for _, _, z, _ in p2p], dtype=float).reshape(-1, 3)
    g.p2p_sigma = np.array([s for _, _, _, s in p2p], dtype=float)
    g.p2p_huber = np.full(len(p2p), float(huber_k))
    g.p2p_order = np.array([s for s, _, _, _ in p2p], dtype=np.int32)
    g.ter_prev = np.array([look(qidx, k[0], "point") for _, k, _, _ in ter], dtype=np.int32)
    g.ter_cur = np.array([look(qidx, k[1], "point") for _, k, _, _ in ter], dtype=np.int32)
    g.ter_motion = np.array([look(pidx, k[2], "pose") for _, k, _, _ in ter], dtype=np.int32)
    g.ter_sigma = np.array([s for _, _, _, s in ter], dtype=float)
    g.ter_huber = np.full(len(ter), float(huber_k))
    g.ter_order = np.array([s for s, _, _, _ in ter], dtype=np.int32)
    g.btw_a = np.array([look(pidx, k[0], "pose") for _, k, _, _ in btw], dtype=np.int32)
    g.btw_b = np.array([look(pidx, k[1], "pose") for _, k, _, _ in btw], dtype=np.int32)
    g.btw_z = np.array([z for _, _, z, _ in btw], dtype=float).reshape(-1, 12)
    g.btw_sigmas = np.array([s for _, _, _, s in btw], dtype=float).reshape(-1, 6)
    g.btw_order = np.array([s for s, _, _, _ in btw], dtype=np.int32)
    g.pri_a = np.array([look(pidx, k[0], "pose") for _, k, _, _ in pri], dtype=np.int32)
    g.pri_z = np.array([z for _, _, z, _ in pri], dtype=float).reshape(-1, 12)
    g.pri_sigmas = np.array([s for _, _, _, s in pri], dtype=float).reshape(-1, 6)
    g.pri_order = np.array([s for s, _, _, _ in pri], dtype=np.int32)
    return g
