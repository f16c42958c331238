import sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
import dynosam_b200 as d
which = sys.argv[1:] or ["cfg2", "cfg3"]
CFG = {"cfg2": dict(), "cfg3": dict(n_frames=1000, n_objects=20, static_per_frame=900, dynamic_per_object=100, object_span=300, seed=1003),
       "mid": dict(n_frames=120, n_objects=4, static_per_frame=120, dynamic_per_object=25, seed=11)}
for name in which:
    kw = CFG[name]
    t = time.time(); s = d.Scene.synthetic(**kw); g = s.full_batch_graph(); print(name, "build", time.time() - t, g.n_pose, g.n_point, g.n_p2p, g.n_ternary, g.n_between, flush=True)
    p = d.default_lm_params(); p.max_iterations = 5; p.relative_error_tol = 0; p.absolute_error_tol = 0; p.verbose = int(__import__("os").environ.get("DV", "2"))
    t = time.time(); lm = d.GpuLevenbergMarquardt(g, p); print("load", time.time() - t, flush=True)
    t = time.time(); lm.optimize(); print("opt wall", time.time() - t); print(lm.timing(), lm.error_trace, flush=True)
    if name == "mid":
        from oracle_lib import OracleGraph
        ora = OracleGraph.from_graph(g)
        r = ora.optimize(max_iterations=5, relative_error_tol=0.0, absolute_error_tol=0.0)
        print("oracle", r["error_trace"], r["iterations"], r["inner_iterations"])
