"""One rank of the 2-GPU parity check of the sharded solve (launched by test_gpu_sharded.py under torchrun).
Rank 0 also solves the whole graph on its own GPU; both ranks then solve it sharded and compare."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import dynosam_b200 as d
from dynosam_b200 import sharded
import scenes


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dist.init_process_group("gloo")
    torch.cuda.set_device(local)
    report = {}
    for name, scene in (("cfg1", scenes.simulator_cfg1()), ("medium", scenes.medium_noisy())):
        g = scene.full_batch_graph()
        p = d.default_lm_params()
        p.device = local
        p.max_iterations = 6
        p.linear_solver = 4 if name == "medium" else 3   # the sharded exchange with the two-front and with the single-front factorisation
        uid = [sharded.make_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        lm = sharded.ShardedLevenbergMarquardt(g, rank, world, uid[0], p)
        e0 = lm.error()
        ok, dp, dpt, _ = lm.solve_once(1e-3)
        pose, pts = lm.optimize()
        t = lm.timing()
        shares = [None] * world
        dist.all_gather_object(shares, (lm.share.point_global, pts, dpt))
        if rank == 0:
            one = d.GpuLevenbergMarquardt(g, p)
            r_e0 = one.error()
            r_ok, r_dp, r_dpt, _ = one.solve_once(1e-3)
            r_pose, r_pts = one.optimize()
            merged = sharded.merge_points(g.n_point, [(s[0], s[1]) for s in shares])
            merged_d = sharded.merge_points(g.n_point, [(s[0], s[2]) for s in shares])
            tr, rtr = np.array(lm.error_trace), np.array(one.error_trace)
            report[name] = {
                "e0_rel": abs(e0 - r_e0) / max(abs(r_e0), 1e-300), "solved": [bool(ok), bool(r_ok)],
                "step_pose": float(np.abs(dp - r_dp).max() / max(np.abs(r_dp).max(), 1e-300)),
                "step_point": float(np.abs(merged_d - r_dpt).max() / max(np.abs(r_dpt).max(), 1e-300)),
                "iterations": [lm.iterations(), one.iterations()], "inner": [lm.getInnerIterations(), one.getInnerIterations()],
                "trace_rel": float(np.max(np.abs(tr - rtr) / np.maximum(np.abs(rtr), 1e-12))) if len(tr) == len(rtr) else None,
                "pose": float(np.abs(pose - r_pose).max()), "points": float(np.abs(merged - r_pts).max()),
                "collectives": t["collectives"], "collective_bytes": t["collective_bytes"], "band_allreduce_ms": t["band_allreduce_ms"],
                "band_fronts": t["band_fronts"], "single_front_repeats": t["band_single_front_repeats"],
            }
        del lm
        dist.barrier()
    if rank == 0:
        print("SHARDED_REPORT " + json.dumps(report), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
