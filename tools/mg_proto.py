"""CPU prototype (numpy) of the multigrid-preconditioned CG used to pick the V-cycle design before
writing CUDA.  Not product code, not the oracle: a design notebook.  Cell types: 0 solid, 1 fluid, 2 air."""
import sys, time
import numpy as np
SOLID, FLUID, AIR = 0, 1, 2


def pad(a, v=0):
    return np.pad(a, 1, constant_values=v)


class Level:
    def __init__(self, typ, h2inv):
        self.typ = typ                      # (nz,ny,nx) uint8
        self.fl = typ == FLUID
        tp = pad(typ, SOLID)
        nonsolid = tp != SOLID
        self.cnt = (nonsolid[2:, 1:-1, 1:-1].astype(np.int32) + nonsolid[:-2, 1:-1, 1:-1] + nonsolid[1:-1, 2:, 1:-1]
                    + nonsolid[1:-1, :-2, 1:-1] + nonsolid[1:-1, 1:-1, 2:] + nonsolid[1:-1, 1:-1, :-2])
        self.h2inv = h2inv
        self.diag = self.cnt * h2inv
        self.active = self.fl & (self.cnt > 0)
        self.invd = np.where(self.active, 1.0 / np.maximum(self.diag, 1e-300), 0.0)

    def A(self, x):
        xp = pad(x)
        s = xp[2:, 1:-1, 1:-1] + xp[:-2, 1:-1, 1:-1] + xp[1:-1, 2:, 1:-1] + xp[1:-1, :-2, 1:-1] + xp[1:-1, 1:-1, 2:] + xp[1:-1, 1:-1, :-2]
        return np.where(self.active, self.diag * x - self.h2inv * s, 0.0)


def coarsen_types(typ, rule):
    nz, ny, nx = typ.shape
    cz, cy, cx = (nz + 1) // 2, (ny + 1) // 2, (nx + 1) // 2
    t = np.full((2 * cz, 2 * cy, 2 * cx), SOLID, dtype=np.uint8)
    t[:nz, :ny, :nx] = typ
    blocks = t.reshape(cz, 2, cy, 2, cx, 2).transpose(0, 2, 4, 1, 3, 5).reshape(cz, cy, cx, 8)
    anyair = (blocks == AIR).any(-1); anyfl = (blocks == FLUID).any(-1)
    if rule == "air":      # McAdams: Dirichlet wins
        return np.where(anyair, AIR, np.where(anyfl, FLUID, SOLID)).astype(np.uint8)
    else:                  # fluid wins
        return np.where(anyfl, FLUID, np.where(anyair, AIR, SOLID)).astype(np.uint8)


W1 = np.array([1, 3, 3, 1]) / 8.0


def restrict(rf, shape_c):
    nz, ny, nx = rf.shape
    cz, cy, cx = shape_c
    # pad fine to (2c+2) with one extra on each side
    a = np.zeros((2 * cz + 2, 2 * cy + 2, 2 * cx + 2))
    a[1:nz + 1, 1:ny + 1, 1:nx + 1] = rf
    # separable: along x
    def r1(a, axis):
        sl = [slice(None)] * 3
        out = 0
        n = (a.shape[axis] - 2) // 2
        for o, w in enumerate(W1):
            sl[axis] = slice(o, o + 2 * n, 2)
            out = out + w * a[tuple(sl)]
        return out
    a = r1(a, 2); a = r1(a, 1); a = r1(a, 0)
    return a


def prolong(ec, shape_f):
    nz, ny, nx = shape_f
    cz, cy, cx = ec.shape
    def p1(e, axis, nf):
        e = np.moveaxis(e, axis, 0)
        c = e.shape[0]
        ep = np.concatenate([np.zeros((1,) + e.shape[1:]), e, np.zeros((1,) + e.shape[1:])], 0)
        out = np.zeros((2 * c,) + e.shape[1:])
        out[0::2] = 0.75 * ep[1:-1] + 0.25 * ep[:-2]
        out[1::2] = 0.75 * ep[1:-1] + 0.25 * ep[2:]
        return np.moveaxis(out[:nf], 0, axis)
    e = p1(ec, 2, nx); e = p1(e, 1, ny); e = p1(e, 0, nz)
    return e


class MG:
    def __init__(self, typ, h2inv=1.0, rule="air", nu=2, omega=2.0 / 3.0, smoother="jacobi", min_n=4, coarse_sweeps=20, bsmooth=0):
        self.levels = [Level(typ, h2inv)]
        self.nu, self.omega, self.smoother, self.coarse_sweeps, self.bsmooth = nu, omega, smoother, coarse_sweeps, bsmooth
        while min(self.levels[-1].typ.shape) > min_n and self.levels[-1].active.sum() > 0:
            t = coarsen_types(self.levels[-1].typ, rule)
            self.levels.append(Level(t, self.levels[-1].h2inv / 4.0))
        # colour masks
        for L in self.levels:
            z, y, x = np.indices(L.typ.shape)
            L.red = ((x + y + z) & 1) == 0

    def smooth(self, L, x, b, reverse=False, sweeps=None):
        n = self.nu if sweeps is None else sweeps
        for _ in range(n):
            if self.smoother == "jacobi":
                x = x + self.omega * L.invd * (b - L.A(x))
            else:
                order = [~L.red, L.red] if reverse else [L.red, ~L.red]
                for m in order:
                    r = b - L.A(x)
                    x = x + np.where(m, L.invd * r, 0.0)
        return x

    def vcycle(self, l, b):
        L = self.levels[l]
        if l == len(self.levels) - 1:
            x = np.zeros_like(b)
            x = self.smooth(L, x, b, False, self.coarse_sweeps)
            x = self.smooth(L, x, b, True, self.coarse_sweeps)
            return x
        x = np.zeros_like(b)
        x = self.smooth(L, x, b, False)
        r = b - L.A(x)
        Lc = self.levels[l + 1]
        rc = np.where(Lc.active, restrict(np.where(L.active, r, 0.0), Lc.typ.shape), 0.0)
        ec = self.vcycle(l + 1, rc)
        x = x + np.where(L.active, prolong(ec, L.typ.shape), 0.0)
        x = self.smooth(L, x, b, True)
        return x


def pcg(L, b, precond, tol=2.220446049250313e-16, maxit=5000, verbose=False):
    x = np.zeros_like(b); r = b.copy()
    rhs2 = (b * b).sum(); thr = tol * tol * rhs2
    z = precond(r); p = z.copy(); absNew = (r * z).sum()
    hist = []
    for i in range(maxit):
        Ap = L.A(p)
        alpha = absNew / (p * Ap).sum()
        x += alpha * p; r -= alpha * Ap
        rr = (r * r).sum(); hist.append(np.sqrt(rr / rhs2))
        if rr < thr:
            return x, i, hist
        z = precond(r)
        absOld = absNew; absNew = (r * z).sum()
        p = z + (absNew / absOld) * p
    return x, maxit, hist


def problem_from_oracle(o, O):
    C = o.fgrid(O.F_CONTAINER); solid = o.solid
    typ = np.where(solid == 1, SOLID, np.where(C > 0, FLUID, AIR)).astype(np.uint8)
    b = o.fgrid(O.F_DIVER).astype(np.float64)
    return typ, b


if __name__ == "__main__":
    sys.path.insert(0, "oracle"); sys.path.insert(0, ".")
    import oracle_py as O
    from fluid_simulation_b200 import scenes
    which = sys.argv[1] if len(sys.argv) > 1 else "tank64"
    if which.startswith("tank"):
        n = int(which[4:])
        sc = scenes.dam_break(n=n, fill="full")
        o = O.Oracle(nx=n, ny=n, nz=n, ox=n // 2, oy=n // 2, oz=n // 2, threads=8)
        steps = 0
    elif which.startswith("col"):
        n = int(which[3:])
        sc = scenes.dam_break(n=n, fill="column")
        o = O.Oracle(nx=n, ny=n, nz=n, ox=n // 2, oy=n // 2, oz=n // 2, threads=8, solver=1)
        steps = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    else:
        sc = scenes.cube_drop_default()
        o = O.Oracle(threads=8)
        steps = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    o.set_particles(sc.pos, sc.vel)
    dt = 0.1
    for s in range(steps):
        dt, st = o.step(dt)
        print("advance", s, st.passes, st.pcg_iters_total, flush=True)
    o.clear_vel(); o.p2g(); o.classify(); o.rhs_div(dt)
    typ, b = problem_from_oracle(o, O)
    if (b == 0).all():
        rng = np.random.default_rng(0); b = np.where(typ == FLUID, rng.normal(size=typ.shape), 0.0)
    print("n_fluid", (typ == FLUID).sum(), "shape", typ.shape)
    L = Level(typ, 1.0)
    b = np.where(L.active, b, 0.0)
    t = time.time(); x, it, h = pcg(L, b, lambda r: L.invd * r); print("jacobi-pcg iters", it, "%.1fs" % (time.time() - t), flush=True)
    for cfg in [dict(rule="air", nu=2, smoother="jacobi"), dict(rule="air", nu=1, smoother="rbgs"), dict(rule="air", nu=2, smoother="rbgs"),
                dict(rule="fluid", nu=2, smoother="rbgs")]:
        mg = MG(typ, **cfg)
        t = time.time(); x2, it2, h2 = pcg(L, b, lambda r: mg.vcycle(0, r), maxit=300)
        print(cfg, "levels", len(mg.levels), "iters", it2, "err vs jacobi %.2e" % (np.linalg.norm(x2 - x) / np.linalg.norm(x)),
              "rate %.3f" % (h2[-1] / h2[0]) ** (1.0 / max(1, len(h2) - 1)), "%.1fs" % (time.time() - t), flush=True)
