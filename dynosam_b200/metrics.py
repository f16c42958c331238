"""Post-solve accuracy metrics against a scene's ground truth (host side, numpy).

pose_error / relative_pose_error follow the reference's TRErrorPair (src/utils/Metrics.cc:34-55): E = M^-1 M_hat,
translation error |t_E|, rotation error = angle of R_E. Camera ATE / RPE are their RMSE over frames; the object
motion error (AME in dynosam_utils/evaluation/core/metrics.py:49-54) is the same error on H_gt^-1 H_est."""
import numpy as np

from .api import key_decode


def _split(Rt):
    Rt = np.asarray(Rt, dtype=float).reshape(-1, 12)
    return Rt[:, :9].reshape(-1, 3, 3), Rt[:, 9:12]


def pose_error(M, M_hat):
    """(translation error [m], rotation error [rad]) per pose pair; inputs (n, 12) row-major R | t."""
    R, t = _split(M)
    Rh, th = _split(M_hat)
    Rt_ = np.transpose(R, (0, 2, 1))
    tE = np.einsum("nij,nj->ni", Rt_, th - t)
    RE = np.einsum("nij,njk->nik", Rt_, Rh)
    cos = np.clip((np.trace(RE, axis1=1, axis2=2) - 1.0) * 0.5, -1.0, 1.0)
    return np.linalg.norm(tE, axis=1), np.arccos(cos)


def _compose_inv_a_b(A, B):
    """A^-1 B for stacks of (n, 12) poses"""
    Ra, ta = _split(A)
    Rb, tb = _split(B)
    Rat = np.transpose(Ra, (0, 2, 1))
    R = np.einsum("nij,njk->nik", Rat, Rb)
    t = np.einsum("nij,nj->ni", Rat, tb - ta)
    return np.concatenate([R.reshape(-1, 9), t], axis=1)


def relative_pose_error(M_ref, M_cur, M_hat_ref, M_hat_cur):
    return pose_error(_compose_inv_a_b(M_ref, M_cur), _compose_inv_a_b(M_hat_ref, M_hat_cur))


def rmse(x):
    x = np.asarray(x, dtype=float)
    return float(np.sqrt(np.mean(x * x))) if x.size else 0.0


def evaluate(scene, pose_keys, pose_Rt):
    """Camera ATE / RPE and object-motion error of the pose variables (keys + (n, 12) values) against scene ground truth."""
    pose_Rt = np.asarray(pose_Rt, dtype=float).reshape(-1, 12)
    gt_cam = scene.gt_camera()
    objs, frames, motions = scene.gt_motions()
    gt_mot = {(int(o), int(f)): m for o, f, m in zip(objs, frames, motions)}
    cam_idx, cam_frame, mot_idx, mot_gt = [], [], [], []
    for i, k in enumerate(pose_keys):
        c, obj, frame, _ = key_decode(int(k))
        if c == "X":
            cam_idx.append(i); cam_frame.append(frame)
        elif c == "H" and (obj, frame) in gt_mot:
            mot_idx.append(i); mot_gt.append(gt_mot[(obj, frame)])
    out = {}
    if cam_idx:
        order = np.argsort(cam_frame)
        idx, fr = np.array(cam_idx)[order], np.array(cam_frame)[order]
        t_e, r_e = pose_error(gt_cam[fr], pose_Rt[idx])
        out["camera_ate_m"], out["camera_ate_rad"] = rmse(t_e), rmse(r_e)
        if len(idx) > 1:
            t_r, r_r = relative_pose_error(gt_cam[fr[:-1]], gt_cam[fr[1:]], pose_Rt[idx[:-1]], pose_Rt[idx[1:]])
            out["camera_rpe_m"], out["camera_rpe_rad"] = rmse(t_r), rmse(r_r)
    if mot_idx:
        t_e, r_e = pose_error(np.array(mot_gt), pose_Rt[np.array(mot_idx)])
        out["object_motion_error_m"], out["object_motion_error_rad"] = rmse(t_e), rmse(r_e)
        out["n_object_motions"] = len(mot_idx)
    return out
