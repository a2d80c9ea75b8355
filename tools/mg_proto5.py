"""smoother study: per-sweep Jacobi weights (Chebyshev-type) instead of a constant omega = 0.8.
pre-sweeps use (w1, w2), post-sweeps the reversed order (w2, w1) so that the V-cycle stays symmetric."""
import sys, time
import numpy as np
sys.path.insert(0, "tools")
from mg_proto import *
from mg_proto2 import MG2
from mg_proto4 import MGc


class MGw(MGc):
    def __init__(self, typ, weights=(0.8, 0.8), **kw):
        self.weights = weights
        super().__init__(typ, **kw)

    def smooth(self, L, x, b, reverse=False, sweeps=None):
        if sweeps is not None:
            return super().smooth(L, x, b, reverse, sweeps)
        ws = self.weights[::-1] if reverse else self.weights
        for w in ws:
            x = x + w * L.invd * (b - L.A(x))
        return x


if __name__ == "__main__":
    sys.path.insert(0, "oracle"); sys.path.insert(0, ".")
    import oracle_py as O
    from fluid_simulation_b200 import scenes
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    sc = scenes.dam_break(n=n, fill="full")
    o = O.Oracle(nx=n, ny=n, nz=n, ox=n // 2, oy=n // 2, oz=n // 2, threads=8)
    o.set_particles(sc.pos, sc.vel)
    o.clear_vel(); o.p2g(); o.classify(); o.rhs_div(0.1)
    typ, b = problem_from_oracle(o, O)
    L = Level(typ, 1.0); b = np.where(L.active, b, 0.0)
    def cheb(a, bb, k):
        c, h = (a + bb) / 2, (bb - a) / 2
        return tuple(1.0 / (c + h * np.cos(np.pi * (2 * i + 1) / (2 * k))) for i in range(k))
    cands = [(0.8, 0.8), cheb(2 / 3, 2, 2), cheb(0.5, 2, 2), cheb(0.4, 2, 2), cheb(1.0, 2, 2), cheb(0.5, 2, 3), cheb(0.3, 2, 3), (0.8, 0.8, 0.8),
             cheb(0.5, 1.9, 2), cheb(0.6, 2, 2)[::-1]]
    for w in cands:
        mg = MGw(typ, weights=w, normalize=True, nu=2, smoother="jacobi", omega=0.8, min_n=16, coarse="cheb", lmin=0.005)
        t = time.time(); x2, it2, h2 = pcg(L, b, lambda r: mg.vcycle(0, r), maxit=200)
        print("weights", tuple(round(x, 3) for x in w), "iters", it2, "%.1fs" % (time.time() - t), flush=True)
