"""Dev script: phase times of one sliding-window solve (cfg4 shape)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import dynosam_b200 as d
scene = d.Scene.synthetic(**bench.WORKLOADS["cfg3"])
p = d.default_lm_params(); p.verbose = int(os.environ.get("DV", "0"))
lm = d.GpuLevenbergMarquardt(None, p)
for (a, b) in [(300, 325), (303, 328), (306, 331)]:
    t0 = time.perf_counter(); g = scene.window_graph(a, b, True); t1 = time.perf_counter()
    lm.load_graph(g); t2 = time.perf_counter()
    lm.optimize(); t3 = time.perf_counter()
    t = lm.timing()
    print("window %d-%d: factors %d poses %d; build %.1f ms, load %.1f ms, optimize %.1f ms (device %.1f), iters %d inner %d" % (a, b, g.n_factors, g.n_pose, 1e3*(t1-t0), 1e3*(t2-t1), 1e3*(t3-t2), t["optimize_ms"], lm.iterations(), lm.getInnerIterations()))
    print("   ", {k: round(v, 2) for k, v in t.items() if k.endswith("_ms") or k in ("kernel_launches", "n_linearize", "n_solves", "n_error_passes", "half_bandwidth", "reduced_dim")})
