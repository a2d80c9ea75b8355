"""coarsest-level solver study: 16^3 coarsest grid with 41 Jacobi sweeps vs Chebyshev(40) vs deeper hierarchy"""
import sys, time
import numpy as np
sys.path.insert(0, "tools")
from mg_proto import *
from mg_proto2 import MG2


class MGc(MG2):
    def __init__(self, typ, coarse="jacobi", lmin=0.005, **kw):
        self.coarse = coarse; self.lmin = lmin
        super().__init__(typ, **kw)

    def vcycle(self, l, b):
        L = self.levels[l]
        if l == len(self.levels) - 1:
            if self.coarse == "jacobi":
                x = np.zeros_like(b)
                for _ in range(41):
                    x = x + self.omega * L.invd * (b - L.A(x))
                return x
            # Chebyshev on D^-1 A, eigenvalues in [lmin, 2]
            lmax, lmin = 2.0, self.lmin
            theta, delta = (lmax + lmin) / 2, (lmax - lmin) / 2
            sigma = theta / delta; rho = 1.0 / sigma
            x = np.zeros_like(b); r = L.invd * b; d = r / theta
            for _ in range(40):
                x = x + d
                r = r - L.invd * L.A(d)
                rho_new = 1.0 / (2 * sigma - rho)
                d = rho_new * rho * d + (2 * rho_new / delta) * r
                rho = rho_new
            return x
        return super().vcycle(l, b)


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
    for cfg in [dict(min_n=16, coarse="jacobi"), dict(min_n=16, coarse="cheb", lmin=0.005), dict(min_n=16, coarse="cheb", lmin=0.02),
                dict(min_n=4, coarse="jacobi")]:
        mg = MGc(typ, normalize=True, nu=2, smoother="jacobi", omega=0.8, **cfg)
        t = time.time(); x2, it2, h2 = pcg(L, b, lambda r: mg.vcycle(0, r), maxit=300)
        print(cfg, "levels", len(mg.levels), "coarsest", mg.levels[-1].typ.shape, "iters", it2, "%.1fs" % (time.time() - t), flush=True)
