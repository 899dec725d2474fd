"""A/B of the fused sweep on wide rows: python tools/ab_fused.py <package dir> [nvoxel rows iterations] -- the package dir may be a checkout of another commit built beside this one"""
import importlib.util, json, os, sys
import numpy as np
pkg_dir = os.path.abspath(sys.argv[1])
V = int(sys.argv[2]) if len(sys.argv) > 2 else 200000
P = int(sys.argv[3]) if len(sys.argv) > 3 else 25000
I = int(sys.argv[4]) if len(sys.argv) > 4 else 10
spec = importlib.util.spec_from_file_location("sartpkg", os.path.join(pkg_dir, "__init__.py"), submodule_search_locations=[pkg_dir])
sart = importlib.util.module_from_spec(spec); sys.modules["sartpkg"] = sart; spec.loader.exec_module(sart)
with sart.Engine(P, V, device=0, beta_laplace=0.0, max_iterations=I, fixed_iterations=True) as eng:
    eng.fill_synthetic(42)
    eng.finalize()
    rng = np.random.default_rng(1)
    g = eng.forward_project(rng.random(V))
    x, st, it = eng.solve(g)
    res = []
    for k in range(3):
        x, st, it = eng.solve(g, x)
        s = eng.stats()
        res.append(s.iterate_ms / (it + 1))
    out = {"pkg": sys.argv[1], "nvoxel": V, "rows": P, "iterations": it, "sweep_ms": res, "hbm_gbs": P * V * 4 / min(res) / 1e6,
           "sweep_mode": s.sweep_mode, "x_hash": float(np.sum(x * np.arange(1, V + 1)))}
    for k in ("fused_cluster_size", "fused_cluster_count", "tail_ms_avg"):
        if hasattr(s, k): out[k] = getattr(s, k)
print("WIDE " + json.dumps(out))
