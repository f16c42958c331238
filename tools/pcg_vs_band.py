"""Dev script: one damped solve (lambda = 1e-5 at the front-end values) by the matrix-free block-Jacobi PCG and by the band
Cholesky, on BASELINE configs 2 and 3: iterations, residual reached, wall time. Evidence for DESIGN.md (why the direct
solve is the default reduced solver).   python tools/pcg_vs_band.py [cfg2 cfg3]"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
import dynosam_b200 as d
import torch

out = {}
for name in (sys.argv[1:] or ["cfg2", "cfg3"]):
    scene, g = bench.build_graph(bench.WORKLOADS[name])
    res = {}
    for label, solver, cap in (("band_cholesky", 0, 0), ("pcg_2000", 1, 2000), ("pcg_20000", 1, 20000)):
        p = d.default_lm_params()
        p.linear_solver = solver
        if cap:
            p.pcg_max_iterations = cap
        lm = d.GpuLevenbergMarquardt(g, p)
        lm.solve_once(1e-5)           # warm-up (allocations, caches)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        ok, dp, dpt, iters = lm.solve_once(1e-5)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        res[label] = {"solved": ok, "iterations_or_solves": iters, "ms": 1e3 * dt, "residual": lm.solve_residual()["last"], "step_norm": float(np.linalg.norm(dp))}
        del lm
    res["pcg_step_vs_band_step"] = None
    out[name] = res
    print(name, json.dumps(res), flush=True)
json.dump(out, open(os.path.join("gpurun_out", "r2_pcg_vs_band.json"), "w"), indent=1)
