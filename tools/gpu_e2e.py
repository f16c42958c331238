"""Dev script: where does the end-to-end (host buffers -> optimised values) time go?  python gpu_e2e.py [cfg3|cfg2]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
import dynosam_b200 as d
import torch

name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
scene, g = bench.build_graph(bench.WORKLOADS[name])
p = bench.lm_params(0)
p.verbose = int(os.environ.get("DV", "1"))
lm = d.GpuLevenbergMarquardt(None, p)
for rep in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    lm.set_variables(g.pose_keys, g.pose_Rt, g.point_keys, g.point_xyz)
    t1 = time.perf_counter()
    lm.add_pose_to_point(g.p2p_pose, g.p2p_point, g.p2p_z, 0.06, 1e-4)
    lm.add_motion_ternary(g.ter_prev, g.ter_cur, g.ter_motion, 0.01, 1e-4)
    lm.add_between(g.btw_a, g.btw_b, g.btw_z, g.btw_sigmas[0])
    lm.add_prior(g.pri_a, g.pri_z, g.pri_sigmas[0])
    t2 = time.perf_counter()
    lm.finalize()
    t3 = time.perf_counter()
    lm.optimize()
    t4 = time.perf_counter()
    print("rep %d: set_variables %.1f ms, add_* %.1f ms, finalize %.1f ms, optimize+values %.1f ms" % (rep, 1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2), 1e3 * (t4 - t3)))
    t0 = time.perf_counter()
    lm.load_host_arrays(g)
    t1 = time.perf_counter()
    print("   load_host_arrays %.1f ms" % (1e3 * (t1 - t0)))
