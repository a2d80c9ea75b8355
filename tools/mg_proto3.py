"""fp32 V-cycle inside fp64 PCG: does the iteration count survive? (design notebook)"""
import sys, time
import numpy as np
sys.path.insert(0, "tools")
from mg_proto import *
from mg_proto2 import MG2


class MG32(MG2):
    def vcycle32(self, r64, scale):
        r = (r64 * scale).astype(np.float32)
        z = self._v(0, r)
        return z.astype(np.float64) / scale

    def _sm(self, L, x, b, sweeps):
        om = np.float32(self.omega)
        invd = L.invd.astype(np.float32); diag = L.diag.astype(np.float32); h = np.float32(L.h2inv)
        for _ in range(sweeps):
            xp = np.pad(x, 1)
            s = xp[2:, 1:-1, 1:-1] + xp[:-2, 1:-1, 1:-1] + xp[1:-1, 2:, 1:-1] + xp[1:-1, :-2, 1:-1] + xp[1:-1, 1:-1, 2:] + xp[1:-1, 1:-1, :-2]
            Ax = np.where(L.active, diag * x - h * s, np.float32(0))
            x = (x + om * invd * (b - Ax)).astype(np.float32)
        return x

    def _v(self, l, b):
        L = self.levels[l]
        if l == len(self.levels) - 1:
            return self._sm(L, np.zeros_like(b), b, 2 * self.coarse_sweeps)
        x = self._sm(L, np.zeros_like(b), b, self.nu)
        xp = np.pad(x, 1)
        s = xp[2:, 1:-1, 1:-1] + xp[:-2, 1:-1, 1:-1] + xp[1:-1, 2:, 1:-1] + xp[1:-1, :-2, 1:-1] + xp[1:-1, 1:-1, 2:] + xp[1:-1, 1:-1, :-2]
        r = np.where(L.active, b - (L.diag.astype(np.float32) * x - np.float32(L.h2inv) * s), np.float32(0)).astype(np.float32)
        P, fi = self.P[l]
        Lc = self.levels[l + 1]
        P32 = P.astype(np.float32)
        rc = (P32.T @ r[fi[:, 0], fi[:, 1], fi[:, 2]] / np.float32(8)).reshape(Lc.typ.shape).astype(np.float32)
        rc = np.where(Lc.active, rc, np.float32(0))
        ec = self._v(l + 1, rc)
        e = np.zeros_like(b); e[fi[:, 0], fi[:, 1], fi[:, 2]] = P32 @ ec.ravel()
        x = (x + np.where(L.active, e, np.float32(0))).astype(np.float32)
        return self._sm(L, x, b, self.nu)


def pcg_scaled(L, b, mg, tol=2.220446049250313e-16, maxit=300):
    x = np.zeros_like(b); r = b.copy()
    rhs2 = (b * b).sum(); thr = tol * tol * rhs2
    rr = rhs2
    z = mg.vcycle32(r, 1.0 / np.sqrt(rr)); p = z.copy(); absNew = (r * z).sum()
    hist = []
    for i in range(maxit):
        Ap = L.A(p)
        alpha = absNew / (p * Ap).sum()
        x += alpha * p; r -= alpha * Ap
        rr_prev = rr; rr = (r * r).sum(); hist.append(np.sqrt(rr / rhs2))
        if rr < thr:
            return x, i, hist
        z = mg.vcycle32(r, 1.0 / np.sqrt(rr_prev))
        absOld = absNew; absNew = (r * z).sum()
        p = z + (absNew / absOld) * p
    return x, maxit, hist


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
    mg = MG32(typ, normalize=True, nu=2, smoother="jacobi", omega=0.8)
    x64, it64, h64 = pcg(L, b, lambda r: mg.vcycle(0, r), maxit=300)
    x32, it32, h32 = pcg_scaled(L, b, mg)
    print("fp64 V-cycle iters", it64, " fp32 V-cycle iters", it32, " solution diff %.2e" % (np.linalg.norm(x32 - x64) / np.linalg.norm(x64)))
    print("final rel resid fp32-precond: %.2e" % h32[-1], " true resid %.2e" % (np.linalg.norm(b - L.A(x32)) / np.linalg.norm(b)))
