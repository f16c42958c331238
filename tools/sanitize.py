"""compute-sanitizer target: BASELINE config 1 (the reference's simulator scene) and the 120-frame scene with a real band and
two fronts, three LM iterations each, through the C ABI.  Run as
    compute-sanitizer --tool memcheck  python tools/sanitize.py
    compute-sanitizer --tool racecheck python tools/sanitize.py
(summaries are kept under profiles/)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import dynosam_b200 as d  # noqa: E402
import scenes  # noqa: E402

which = sys.argv[1:] or ["cfg1", "medium"]
for name in which:
    g = (scenes.simulator_cfg1() if name == "cfg1" else scenes.medium_noisy()).full_batch_graph()
    p = d.default_lm_params()
    p.max_iterations = 3
    p.relative_error_tol = 0.0
    p.absolute_error_tol = 0.0
    lm = d.GpuLevenbergMarquardt(g, p)
    e0 = lm.error()
    lm.optimize()
    print(name, "factors", g.n_factors, "chi2", e0, "->", lm.error(), "iterations", lm.iterations(), "fronts", lm.timing()["band_fronts"], flush=True)
