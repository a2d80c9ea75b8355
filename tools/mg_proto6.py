"""per-level smoothing schedules: the product runs 2 + 2 damped-Jacobi sweeps with weights (0.6, 1.7) on
every level; on the small levels (64^3 and below at 256^3) every sweep is a latency-bound kernel of
5-8 us, ~100 us of a 608 us iteration.  Which levels can do with 1 + 1 sweeps (first pre-sweep from a zero
guess is free: x1 = w D^-1 b comes out of the producing kernel) without costing CG iterations?
usage: python tools/mg_proto6.py [n]   (n = 128: levels 128, 64, 32, 16 + Chebyshev solve on 8^3)"""
import sys, time
import numpy as np
sys.path.insert(0, "tools")
from mg_proto import *
from mg_proto2 import MG2
from mg_proto4 import MGc


class MGl(MGc):
    """weights_by_level[l] = tuple of Jacobi weights of the pre-sweeps of level l (post = reversed)."""
    def __init__(self, typ, weights_by_level, **kw):
        self.wl = weights_by_level
        super().__init__(typ, **kw)

    def vcycle(self, l, b):
        self._l = l
        return super().vcycle(l, b)

    def smooth(self, L, x, b, reverse=False, sweeps=None):
        if sweeps is not None:
            return super().smooth(L, x, b, reverse, sweeps)
        l = self.levels.index(L)
        w = self.wl[min(l, len(self.wl) - 1)]
        for wk in (w[::-1] if reverse else w):
            x = x + wk * L.invd * (b - L.A(x))
        return x


if __name__ == "__main__":
    sys.path.insert(0, "oracle"); sys.path.insert(0, ".")
    import oracle_py as O
    from fluid_simulation_b200 import scenes
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    sc = scenes.dam_break(n=n, fill="full")
    o = O.Oracle(nx=n, ny=n, nz=n, ox=n // 2, oy=n // 2, oz=n // 2, threads=8)
    o.set_particles(sc.pos, sc.vel)
    o.clear_vel(); o.p2g(); o.classify(); o.rhs_div(0.1)
    typ, b = problem_from_oracle(o, O)
    L = Level(typ, 1.0); b = np.where(L.active, b, 0.0)
    P = (0.6, 1.7)                      # the product's weights
    cands = {
        "product: (0.6,1.7) everywhere": [P],
        "L>=2: 1 sweep w=0.8": [P, P, (0.8,)],
        "L>=2: 1 sweep w=1.0": [P, P, (1.0,)],
        "L>=1: 1 sweep w=0.8": [P, (0.8,)],
        "L>=1: 1 sweep w=1.0": [P, (1.0,)],
        "L>=3: 1 sweep w=0.8": [P, P, P, (0.8,)],
        "L0: 1 sweep w=0.8, rest product": [(0.8,), P],
        "all: 1 sweep w=0.8": [(0.8,)],
        "L>=2: 3 sweeps cheb": [P, P, (0.52, 0.93, 1.87)],
    }
    for name, wl in cands.items():
        mg = MGl(typ, wl, normalize=True, nu=2, smoother="jacobi", omega=0.8, min_n=8, coarse="cheb", lmin=0.005)
        t = time.time(); x2, it2, h2 = pcg(L, b, lambda r: mg.vcycle(0, r), maxit=200)
        print(f"{name:36s} levels {len(mg.levels)} iters {it2 + 1:3d}  last rel {h2[-1]:.2e}  {time.time() - t:.1f}s", flush=True)
