#!/usr/bin/env python
"""Compare the SASS of every kernel of csrc/*.cu between a git revision and the working tree (no GPU
needed: nvcc cross-compiles).  Used when experimental variants are added without GPU time at hand: the
kernels of the default path must come out instruction-for-instruction identical to the last revision
that was measured and parity-tested on the GPU (adding kernels to a translation unit can move the
register allocation of its neighbours -- that is why mg_h16.cu is a file of its own).
usage: python tools/sass_diff.py <git-rev>"""
import collections
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-c"]


def sass(obj):
    out = subprocess.run([os.path.join(os.path.dirname(NVCC), "cuobjdump"), "-sass", obj], capture_output=True, text=True).stdout
    funcs, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1); funcs[cur] = []; continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(.*?);", line)
        if m and cur is not None:
            funcs[cur].append(re.sub(r"\s+", " ", m.group(1)))
    return funcs


def key(mangled):
    d = subprocess.run(["c++filt", mangled], capture_output=True, text=True).stdout.strip()
    d = re.sub(r"\(.*", "", d)                    # drop the parameter list (trailing parameters may be added)
    d = re.sub(r"^void ", "", d)
    return d.replace("<float>", "").replace(", float, float>", ">").replace(", float>", ">")   # defaulted storage types


def compile_tree(csrc, inc, out):
    procs = []
    for f in sorted(os.listdir(csrc)):
        if f.endswith(".cu"):
            o = os.path.join(out, f[:-3] + ".o")
            procs.append((f, o, subprocess.Popen([NVCC] + FLAGS + ["-I", inc, "-I", csrc, os.path.join(csrc, f), "-o", o],
                                                 stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    objs = {}
    for f, o, p in procs:
        log = p.communicate()[0]
        if p.returncode != 0:
            sys.exit(f"nvcc failed on {f}:\n{log}")
        objs[f] = o
    return objs


def main():
    rev = sys.argv[1] if len(sys.argv) > 1 else "HEAD"
    bad = 0
    with tempfile.TemporaryDirectory() as tmp:
        base = os.path.join(tmp, "base"); os.makedirs(base)
        ar = subprocess.run(["git", "-C", ROOT, "archive", rev, "fluid-simulation_b200/csrc", "include"], capture_output=True)
        subprocess.run(["tar", "-x", "-C", base], input=ar.stdout, check=True)
        os.makedirs(os.path.join(tmp, "o")); os.makedirs(os.path.join(tmp, "n"))
        old = compile_tree(os.path.join(base, "fluid-simulation_b200", "csrc"), os.path.join(base, "include"), os.path.join(tmp, "o"))
        new = compile_tree(os.path.join(ROOT, "fluid-simulation_b200", "csrc"), os.path.join(ROOT, "include"), os.path.join(tmp, "n"))
        for f, o in old.items():
            if f not in new:
                print(f"{f}: removed"); continue
            so = {key(k): v for k, v in sass(o).items()}
            sn = {key(k): v for k, v in sass(new[f]).items() if "__half" not in key(k)}
            same = 0
            for k, v in so.items():
                if k not in sn:
                    print(f"{f}: {k}: not in the working tree"); bad += 1
                elif sn[k] != v:
                    print(f"{f}: {k}: DIFFERENT ({len(v)} vs {len(sn[k])} instructions)"); bad += 1
                else:
                    same += 1
            print(f"{f}: {same}/{len(so)} kernels of {rev} identical; {len(sn) - len(set(sn) & set(so))} new")
        for f in new:
            if f not in old:
                print(f"{f}: new file, {len(sass(new[f]))} kernels")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
