"""storage-precision study: the V-cycle's level-0 sweeps move 75 B/node in fp32, ~260 of the
608 us of an iteration at 256^3.  Would fp16 / bf16 STORAGE of the level-0 vectors (arithmetic stays
fp32) keep the CG iteration count?  Emulated here by rounding every stored level-l array (b, xa, xb,
residual) to the storage type, l < qlevels; the other levels are rounded to fp32 like the product.
fp16 needs the O(1)-scaled residual the product already uses (entries ~1/sqrt(N): scaled by sqrt(N)).
usage: python tools/mg_proto7.py [n]"""
import sys, time
import numpy as np
sys.path.insert(0, "tools")
from mg_proto import *
from mg_proto2 import MG2
from mg_proto4 import MGc


def to_bf16(a):
    a32 = np.ascontiguousarray(a, dtype=np.float32)
    u = a32.view(np.uint32).astype(np.uint64)
    u = ((u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000).astype(np.uint32)     # round to nearest even
    return u.view(np.float32).astype(np.float64)


class MGq(MGc):
    def __init__(self, typ, weights=(0.6, 1.7), qtype="f32", qlevels=0, qkinds="bxr", **kw):
        self.weights, self.qtype, self.qlevels, self.qkinds = weights, qtype, qlevels, qkinds
        super().__init__(typ, **kw)

    def q(self, l, a, kind="x"):
        """kind: b = right-hand side, x = solution iterates, r = residual handed to the restriction"""
        if l >= self.qlevels or self.qtype == "f32" or kind not in self.qkinds:
            return a.astype(np.float32).astype(np.float64)
        if self.qtype == "bf16":
            return to_bf16(a)
        s = self.scale
        return (a * s).astype(np.float16).astype(np.float64) / s

    def vcycle(self, l, b):
        L = self.levels[l]
        if l == 0:
            nb = np.linalg.norm(b)
            self.scale = np.sqrt(L.active.sum()) / (nb if nb > 0 else 1.0)
        if l == len(self.levels) - 1:
            return super().vcycle(l, b)
        b = self.q(l, b, "b")
        if l == 0:
            self.last_b = b                 # what the product's last sweep reads: r.z is taken as b.z (DOT = 2 in mg.cu)
        x = np.zeros_like(b)
        for i, w in enumerate(self.weights):      # kinds 1, 2: outputs of the pre-sweeps
            x = self.q(l, x + w * L.invd * (b - L.A(x)), "x" if "x" in self.qkinds else str(i + 1))
        r = self.q(l, b - L.A(x), "r")
        P, fi = self.P[l]
        Lc = self.levels[l + 1]
        rc = (P.T @ r[fi[:, 0], fi[:, 1], fi[:, 2]] / 8.0).reshape(Lc.typ.shape)
        rc = np.where(Lc.active, rc, 0.0)
        ec = self.vcycle(l + 1, rc)
        e = np.zeros_like(b); e[fi[:, 0], fi[:, 1], fi[:, 2]] = P @ ec.ravel()
        x = self.q(l, x + np.where(L.active, e, 0.0), "x" if "x" in self.qkinds else "3")       # kind 3: after the correction
        for i, w in enumerate(self.weights[::-1]):                                             # kinds 4, 5: post-sweeps (5 = z)
            x = self.q(l, x + w * L.invd * (b - L.A(x)), "x" if "x" in self.qkinds else str(4 + i))
        return x


def pcg_flex(L, b, precond, tol=2.220446049250313e-16, maxit=200, flexible=False, rz=None):
    """pcg of mg_proto.py with an optional flexible (Polak-Ribiere) beta = (r_new - r_old).z_new / r_old.z_old"""
    x = np.zeros_like(b); r = b.copy()
    rhs2 = (b * b).sum(); thr = tol * tol * rhs2
    z = precond(r); p = z.copy(); absNew = (r * z).sum() if rz is None else rz(z)
    for i in range(maxit):
        Ap = L.A(p)
        alpha = absNew / (p * Ap).sum()
        x += alpha * p; r_old = r.copy(); r -= alpha * Ap
        rr = (r * r).sum()
        if rr < thr:
            return x, i + 1
        z = precond(r)
        absOld = absNew; absNew = (r * z).sum() if rz is None else rz(z)
        beta = ((r - r_old) * z).sum() / absOld if flexible else absNew / absOld
        p = z + beta * p
    return x, maxit


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
    xref = None
    full = [("f32", 0, False, "bxr"), ("f16", 1, False, "bxr"), ("f16", 1, True, "bxr"), ("bf16", 1, False, "bxr"), ("bf16", 1, True, "bxr"),
            ("f16", 2, False, "bxr"), ("f16", 9, False, "bxr")]
    kinds = [("f32", 0, False, "bxr"), ("f16", 1, "rzb", "br"), ("f16", 1, False, "br"), ("f32", 1, "rzb", "br"), ("f16", 1, False, "xr"), ("f16", 1, False, "b"), ("f16", 1, False, "r"), ("f16", 1, False, "x"),
             ("bf16", 1, False, "r"), ("bf16", 1, True, "xr")]
    parts = [("f32", 0, False, "bxr"), ("f16", 1, "rzb", "br1"), ("f16", 1, "rzb", "br2"), ("f16", 1, "rzb", "br3"), ("f16", 1, "rzb", "br4"),
             ("f16", 1, "rzb", "br5"), ("f16", 1, "rzb", "br1234"), ("f16", 1, "rzb", "br234"), ("f16", 1, "rzb", "br12")]
    sel = sys.argv[2] if len(sys.argv) > 2 else "full"
    for qtype, ql, flex, qk in {"kinds": kinds, "parts": parts}.get(sel, full):
        mg = MGq(typ, qtype=qtype, qlevels=ql, qkinds=qk, normalize=True, nu=2, smoother="jacobi", omega=0.8, min_n=8, coarse="cheb", lmin=0.005)
        t = time.time()
        if flex == "rzb":                 # r.z := b_stored . z, as the product's fused dot does
            x, it = pcg_flex(L, b, lambda r: mg.vcycle(0, r), rz=lambda z: (mg.last_b * z).sum())
        else:
            x, it = pcg_flex(L, b, lambda r: mg.vcycle(0, r), flexible=flex)
        if xref is None:
            xref = x
        print(f"storage {qtype:5s} of {qk:6s} on levels < {ql}  flexible beta {flex!s:5s}  iters {it:3d}  |x - x_f32| / |x| = "
              f"{np.linalg.norm(x - xref) / np.linalg.norm(xref):.1e}  {time.time() - t:.1f}s", flush=True)
