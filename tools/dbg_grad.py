import os, sys, math
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import liverrenderer_b200 as mi
from tests.test_gpu_northstar import _cube_scene, _window_ranges
res, scale, lo, hi = 16, 2.0, 6, 10
w = hi - lo; spp = 1 << 21
def mk(rho, alb, md, se, tracking="reference"):
    d = _cube_scene(16, res, alb, rho, scale, max_depth=md, max_value=1.0, tracking=tracking)
    d["cube"]["interior"]["sample_emitters"] = se
    return mi.load_dict(d)
def window_mean(rho, seed, alb, md, se):
    ctx = mi.Context(0).upload(mk(rho, alb, md, se))
    v = np.mean([ctx.sample_lanes(seed, spp, b, e)[:, 0].astype(np.float64).mean() for b, e in _window_ranges(res, spp, lo, lo, w, w)]); ctx.close(); return v
eps = 0.02
for name, alb, md, se in (("a.8 md2 nee", .8, 2, True), ("a.8 md2 no-nee", .8, 2, False), ("a.4 md2 nee", .4, 2, True), ("a.8 md8 nee", .8, 8, True), ("a.8 md8 no-nee", .8, 8, False)):
    fds = [(window_mean(0.6 + eps, s, alb, md, se) - window_mean(0.6 - eps, s, alb, md, se)) / (2 * eps) for s in (1, 2)]
    ctx = mi.Context(0).upload(mk(0.6, alb, md, se))
    g_in = np.zeros((res, res, 3), np.float32); g_in[lo:hi, lo:hi, 0] = 1.0 / (w * w)
    vals = []
    for r in range(2):
        ctx.grid_grad_zero(0)
        for b, e in _window_ranges(res, spp, lo, lo, w, w): ctx.render_backward(g_in, seed=100 + r, spp=spp, begin=b, end=e)
        vals.append(float(ctx.grid_grad_read(0).astype(np.float64).sum()))
    ctx.close()
    print(f"{name:16s} FD {np.mean(fds):+.5f} (spread {abs(fds[0]-fds[1]):.1e})  adjoint {np.mean(vals):+.5f} (spread {abs(vals[0]-vals[1]):.1e})  diff {np.mean(vals)-np.mean(fds):+.5f}")
