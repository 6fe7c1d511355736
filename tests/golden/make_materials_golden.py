"""Generate tests/golden/materials_golden.json from the REFERENCE's own generated material code.

Run in the build container only (needs /root/reference):  python tests/golden/make_materials_golden.py
It does not copy reference source: it evaluates the expression text of the S-form functions in
/root/reference/src/energies/*.cpp (MATLAB `ccode` output: plain arithmetic, pow/log/exp) at fixed probe points
with Python floats and stores the resulting NUMBERS.  The probes include the reference tests' literal
s = [1.11 1.2 1.3 .2 .3 .4] (tests/test_dsvd.cpp:119-120).
"""
import json
import math
import os
import re
import numpy as np

REF = "/root/reference/src/energies"
FILES = {"Stable-Neohookean": ("stable_neohookean.cpp", "StableNeohookean"),
         "Neohookean": ("neohookean.cpp", "Neohookean"),
         "Corot": ("corotational.cpp", "Corotational"),
         "FixedCorot": ("fixed_corotational.cpp", "FixedCorotational"),
         "ARAP": ("arap.cpp", "ARAP"),
         "Fung": ("fung.cpp", "Fung")}


def function_body(src, cls, fn, argtype):
    pat = re.compile(r"%s::%s\(const\s+(?:Eigen::)?%s&\s*\w+[^)]*\)\s*\{" % (cls, fn, argtype))
    m = pat.search(src)
    if not m:
        return None
    i = m.end()
    depth = 1
    j = i
    while depth:
        ch = src[j]
        depth += ch == "{"
        depth -= ch == "}"
        j += 1
    return src[i:j - 1]


def evaluate(body, n, s, mu, la):
    env = {"pow": math.pow, "log": math.log, "exp": math.exp, "mu": mu, "la": la, "c": 4.0}
    for k in range(n):
        env["S%d" % (k + 1)] = float(s[k])
    stmts = [t.strip() for t in body.replace("\n", " ").split(";")]
    ret, g, H = None, {}, {}
    zero = False
    for st in stmts:
        if st.startswith("return") and not re.match(r"return\s+[gH]\s*$", st):
            expr = st[len("return"):].strip()
            if "asDiagonal" in expr:   # ARAP::hessian: tmp.asDiagonal() * mu with tmp = 1,1,1,2,2,2
                for k, v in enumerate([1, 1, 1, 2, 2, 2]):
                    H[(k, k)] = v * mu
                zero = True
            else:
                ret = eval(expr, env)
        m = re.match(r"g\((\d)(?:,0)?\)\s*=\s*(.*)$", st)
        if m:
            g[int(m.group(1))] = eval(m.group(2), env)
        m = re.match(r"H\((\d),(\d)\)\s*=\s*(.*)$", st)
        if m:
            H[(int(m.group(1)), int(m.group(2)))] = eval(m.group(3), env)
        if "setZero" in st:
            zero = True
    return ret, g, H, zero


def main():
    rng = np.random.default_rng(7)
    probes6 = [[1.11, 1.2, 1.3, 0.2, 0.3, 0.4], [1, 1, 1, 0, 0, 0], [0.8, 1.3, 0.95, -0.1, 0.05, 0.2]]
    probes6 += (np.array([1, 1, 1, 0, 0, 0]) + rng.uniform(-0.3, 0.3, size=(5, 6))).tolist()
    probes3 = [[1.11, 1.2, 0.2], [1, 1, 0], [0.8, 1.3, -0.1]]
    probes3 += (np.array([1, 1, 0]) + rng.uniform(-0.3, 0.3, size=(5, 3))).tolist()
    params = [(1e6 / (2 * 1.45), (1e6 * 0.45) / (1.45 * 0.1)), (3.0, 7.0)]
    out = {"note": "values computed from the reference's expression text by make_materials_golden.py", "cases": []}
    for name, (fname, cls) in FILES.items():
        src = open(os.path.join(REF, fname)).read()
        for n, argt, matt, probes in ((6, "Vector6d", "Matrix6d", probes6), (3, "Vector3d", "Matrix3d", probes3)):
            be = function_body(src, cls, "energy", argt)
            bg = function_body(src, cls, "gradient", argt)
            bh = function_body(src, cls, "hessian", argt)
            if be is None or bg is None or bh is None:
                continue
            for (mu, la) in params:
                for s in probes:
                    e, _, _, _ = evaluate(be, n, s, mu, la)
                    _, g, _, _ = evaluate(bg, n, s, mu, la)
                    _, _, H, zero = evaluate(bh, n, s, mu, la)
                    Hm = [[H.get((i, j), 0.0 if zero else None) for j in range(n)] for i in range(n)]
                    assert all(v is not None for row in Hm for v in row), (name, n)
                    out["cases"].append({"model": name, "n": n, "mu": mu, "la": la, "s": list(map(float, s)),
                                         "energy": e, "gradient": [g[i] for i in range(n)], "hessian": Hm})
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "materials_golden.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=0)
    print("wrote", path, len(out["cases"]), "cases")


if __name__ == "__main__":
    main()
