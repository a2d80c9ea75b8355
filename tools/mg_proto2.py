"""Variant study: prolongation renormalised at Neumann (solid) coarse neighbours, explicit sparse P."""
import sys, time
import numpy as np, scipy.sparse as sp
sys.path.insert(0, "tools")
from mg_proto import *


def build_P(typ_f, typ_c, normalize):
    nz, ny, nx = typ_f.shape; cz, cy, cx = typ_c.shape
    fi = np.argwhere(typ_f == FLUID)          # (k,j,i)
    rows, cols, vals = [], [], []
    idf = np.arange(fi.shape[0])
    wsum = np.zeros(fi.shape[0])
    ent = []
    for dz in (0, 1):
        for dy in (0, 1):
            for dx in (0, 1):
                w = np.ones(fi.shape[0]); cc = []
                for ax, d in zip((0, 1, 2), (dz, dy, dx)):
                    f = fi[:, ax]; par = f // 2
                    other = np.where(f % 2 == 0, par - 1, par + 1)
                    c = par if d == 0 else other
                    w = w * (0.75 if d == 0 else 0.25)
                    cc.append(c)
                cc = np.stack(cc, 1)
                inb = np.all((cc >= 0) & (cc < np.array(typ_c.shape)), 1)
                tc = np.full(fi.shape[0], SOLID); tc[inb] = typ_c[cc[inb, 0], cc[inb, 1], cc[inb, 2]]
                ent.append((cc, w, tc))
                wsum += np.where(tc != SOLID, w, 0.0)
    for cc, w, tc in ent:
        m = tc == FLUID
        ww = w[m] / wsum[m] if normalize else w[m]
        rows.append(idf[m]); cols.append((cc[m, 0] * cy + cc[m, 1]) * cx + cc[m, 2]); vals.append(ww)
    P = sp.csr_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(fi.shape[0], cz * cy * cx))
    return P, fi


class MG2(MG):
    def __init__(self, typ, normalize=True, **kw):
        super().__init__(typ, **kw)
        self.P = []
        for l in range(len(self.levels) - 1):
            P, fi = build_P(self.levels[l].typ, self.levels[l + 1].typ, normalize)
            self.P.append((P, fi))

    def vcycle(self, l, b):
        L = self.levels[l]
        if l == len(self.levels) - 1:
            x = np.zeros_like(b)
            x = self.smooth(L, x, b, False, self.coarse_sweeps)
            return self.smooth(L, x, b, True, self.coarse_sweeps)
        x = self.smooth(L, np.zeros_like(b), b, False)
        r = b - L.A(x)
        P, fi = self.P[l]
        Lc = self.levels[l + 1]
        rc = (P.T @ r[fi[:, 0], fi[:, 1], fi[:, 2]] / 8.0).reshape(Lc.typ.shape)
        rc = np.where(Lc.active, rc, 0.0)
        ec = self.vcycle(l + 1, rc)
        e = np.zeros_like(b); e[fi[:, 0], fi[:, 1], fi[:, 2]] = P @ ec.ravel()
        x = x + np.where(L.active, e, 0.0)
        return self.smooth(L, x, b, True)


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
    for cfg in [dict(normalize=False, nu=2, smoother="jacobi"), dict(normalize=True, nu=2, smoother="jacobi"),
                dict(normalize=True, nu=2, smoother="rbgs"), dict(normalize=True, nu=1, smoother="rbgs"),
                dict(normalize=True, nu=2, smoother="jacobi", omega=0.8), dict(normalize=True, nu=3, smoother="jacobi", omega=0.8)]:
        mg = MG2(typ, **cfg)
        t = time.time(); x2, it2, h2 = pcg(L, b, lambda r: mg.vcycle(0, r), maxit=300)
        print(cfg, "iters", it2, "rate %.3f" % (h2[-1] / h2[0]) ** (1.0 / max(1, len(h2) - 1)), "%.1fs" % (time.time() - t), flush=True)
