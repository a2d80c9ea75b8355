"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/flipb200.h declares; parameter defaults equal the reference's literals; no CPU fallback."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    txt = open(os.path.join(ROOT, "include", "flipb200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(flipb200_[a-z0-9_]+)\s*\(", txt)))


def test_exports_every_declared_symbol(fb):
    lib = fb.load()
    names = _declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in flipb200.h but not exported"
    assert set(names) == set(fb.EXPORTS)


def test_defaults_are_reference_literals(fb):
    p = fb.default_params()
    assert (p.nx, p.ny, p.nz) == (121, 121, 121) and (p.ox, p.oy, p.oz) == (60, 60, 60)   # fluid.cc:1159
    assert p.wall == 2 and p.relax_div == 10                                                # :1264,1475
    assert (p.gx, p.gy, p.gz) == (0.0, -10.0, 0.0) and p.dx == 1.0 and p.rho == 1.0          # :1357-1358
    assert p.dt_max == 0.1 and p.relax_tol == 0.1 and p.restitution == 0.0                  # :1490,1484,974
    assert p.flip_ratio == 1.0 and p.pcg_tol == np.finfo(np.float64).eps
    assert fb.load().flipb200_version() == 100
    assert fb.load().flipb200_strerror(-3).decode().startswith("PCG breakdown")


def test_no_cpu_fallback(fb):
    """Without a CUDA device create() must fail with FLIPB200_ERR_CUDA, never compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(fb.FlipError) as e:
        fb.Flip(nx=9, ny=9, nz=9, ox=4, oy=4, oz=4)
    assert e.value.status == fb.ERR_CUDA


def test_bad_arguments(fb):
    lib = fb.load()
    p = fb.default_params(wall=0)
    h = C.c_void_p()
    assert lib.flipb200_create(C.byref(p), None, C.byref(h)) == fb.ERR_ARG
    p = fb.default_params(flip_ratio=1.5)
    assert lib.flipb200_create(C.byref(p), None, C.byref(h)) == fb.ERR_ARG
    assert lib.flipb200_step(None, None, None) == fb.ERR_ARG


def test_product_does_not_touch_oracle():
    """The product tree must not reference oracle/ (a product path through the oracle voids parity)."""
    pkg = os.path.join(ROOT, "fluid-simulation_b200")
    for dp, _, fs in os.walk(pkg):
        if os.path.basename(dp) == "build":
            continue
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cc", ".h", ".sh")):
                txt = open(os.path.join(dp, f)).read()
                assert "oracle_py" not in txt and "flip_oracle" not in txt, os.path.join(dp, f)


def test_scenes_counts():
    from fluid_simulation_b200 import scenes
    s = scenes.cube_drop_default()
    assert s.num_particles == 689210            # 10 * 41^3, fluid.cc:1176,1349
    assert np.all(np.abs(s.pos) < 20.5 + 1e-9)
    d = scenes.dam_break(n=32, fill="full")
    assert d.num_particles == 28 * 26 * 28 * 8


def test_struct_layout_matches_the_header(fb, tmp_path):
    """The ctypes mirrors of flipb200_params / flipb200_stats against the C compiler's view of include/flipb200.h: size and
    the offset of every field (a field added on one side only would silently shift everything behind it)."""
    import ctypes as C
    import subprocess
    fields_p = [n for n, _ in fb.Params._fields_]
    fields_s = [n for n, _ in fb.Stats._fields_]
    src = tmp_path / "layout.c"
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "flipb200.h"', 'int main(void) {',
             '  printf("P %zu\\n", sizeof(flipb200_params)); printf("S %zu\\n", sizeof(flipb200_stats));']
    lines += [f'  printf("P.{n} %zu\\n", offsetof(flipb200_params, {n}));' for n in fields_p]
    lines += [f'  printf("S.{n} %zu\\n", offsetof(flipb200_stats, {n}));' for n in fields_s]
    lines += ['  return 0; }']
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.check_call(["/usr/bin/gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    out = dict(l.split() for l in subprocess.check_output([str(exe)], text=True).splitlines())
    assert int(out["P"]) == C.sizeof(fb.Params) and int(out["S"]) == C.sizeof(fb.Stats)
    for n in fields_p:
        assert int(out[f"P.{n}"]) == getattr(fb.Params, n).offset, n
    for n in fields_s:
        assert int(out[f"S.{n}"]) == getattr(fb.Stats, n).offset, n
