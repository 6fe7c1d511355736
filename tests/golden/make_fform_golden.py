"""Golden numbers for the F-form material overloads (the non-mixed Stretch variable) from the
REFERENCE'S OWN generated code: the expression text of energy / gradient / hessian(Vector9d | Vector4d) in
/root/reference/src/energies/{stable_neohookean,neohookean,fixed_corotational}.cpp is evaluated with Python floats at fixed
probe deformation gradients and the NUMBERS are stored (no source is copied).  What is stored per case:
  g_assign / H_assign ... entries written with '='   (Stable-Neohookean, Neohookean: the whole gradient / Hessian BEFORE
                          sim::simple_psd_fix, which is the last statement of those functions)
  g_add / H_add ......... entries written with '+='  (FixedCorot 3D: the volume term added to mu (F - R) / mu (I - dR/dF),
                          which come from iARAP::rotation and are not expression text)
Run in the build container only (needs /root/reference):  python tests/golden/make_fform_golden.py"""
import json
import math
import os
import re

import numpy as np

REF = "/root/reference/src/energies"
CASES = [("Stable-Neohookean", "stable_neohookean.cpp", "StableNeohookean", 9), ("Stable-Neohookean", "stable_neohookean.cpp", "StableNeohookean", 4),
         ("Neohookean", "neohookean.cpp", "Neohookean", 9), ("FixedCorot", "fixed_corotational.cpp", "FixedCorotational", 9)]


def function_body(src, cls, fn, n):
    m = re.search(r"%s::%s\(const\s+(?:Eigen::)?Vector%dd&\s*\w+[^)]*\)\s*\{" % (cls, fn, n), src)
    if not m:
        return None
    i = j = m.end()
    depth = 1
    while depth:
        depth += src[j] == "{"
        depth -= src[j] == "}"
        j += 1
    return src[i:j - 1]


def evaluate(body, F, mu, la):
    env = {"pow": math.pow, "log": math.log, "exp": math.exp, "mu": mu, "la": la}
    out = {"ret": None, "g_assign": {}, "g_add": {}, "H_assign": {}, "H_add": {}}
    body = re.sub(r"//[^\n]*", "", body)                     # comments would glue onto the statement after them
    for st in [t.strip() for t in body.replace("\n", " ").split(";")]:
        m = re.match(r"double\s+(F\d_\d)\s*=\s*F\((\d)\)$", st)
        if m:
            env[m.group(1)] = float(F[int(m.group(2))])
            continue
        m = re.match(r"return\s+(.*)$", st)
        if m and not re.match(r"^[gH]$", m.group(1).strip()):
            try:
                out["ret"] = eval(m.group(1), env)
            except Exception:
                pass
            continue
        m = re.match(r"g\((\d)(?:,0)?\)\s*(\+?=)\s*(.*)$", st)
        if m:
            out["g_add" if m.group(2) == "+=" else "g_assign"][int(m.group(1))] = eval(m.group(3), env)
            continue
        m = re.match(r"H\((\d),(\d)\)\s*(\+?=)\s*(.*)$", st)
        if m:
            out["H_add" if m.group(3) == "+=" else "H_assign"][(int(m.group(1)), int(m.group(2)))] = eval(m.group(4), env)
    return out


def main():
    rng = np.random.default_rng(11)
    probes = {9: [np.array([1, .1, .3, .1, 2, .4, .2, .4, .5])],       # tests/test_dsvd.cpp:76-79, column-major
              4: [np.array([1.1, .1, -.2, .9])]}
    for n in (9, 4):
        d = 3 if n == 9 else 2
        for _ in range(5):
            probes[n].append((np.eye(d) + 0.25 * rng.standard_normal((d, d))).T.reshape(-1))
    params = [(1e6 / (2 * 1.45), (1e6 * 0.45) / (1.45 * 0.1)), (3.0, 7.0)]
    cases = []
    for name, fname, cls, n in CASES:
        src = open(os.path.join(REF, fname)).read()
        be, bg, bh = (function_body(src, cls, f, n) for f in ("energy", "gradient", "hessian"))
        assert bg is not None and bh is not None, (name, n)
        for mu, la in params:
            for F in probes[n]:
                if np.linalg.det(F.reshape(int(n ** .5), -1)) <= 0.05:
                    continue
                e = evaluate(be, F, mu, la)["ret"] if be else None
                g = evaluate(bg, F, mu, la)
                H = evaluate(bh, F, mu, la)
                rec = {"model": name, "n": n, "mu": mu, "la": la, "F": F.tolist(), "energy": e}
                for key, src_d in (("g_assign", g["g_assign"]), ("g_add", g["g_add"])):
                    if src_d:
                        assert len(src_d) == n
                        rec[key] = [src_d[i] for i in range(n)]
                for key, src_d in (("H_assign", H["H_assign"]), ("H_add", H["H_add"])):
                    if src_d:
                        assert len(src_d) == n * n, (name, key, len(src_d))
                        rec[key] = [[src_d[(i, j)] for j in range(n)] for i in range(n)]
                cases.append(rec)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "fform_golden.json")
    json.dump({"note": "values computed from the reference's expression text by make_fform_golden.py", "cases": cases}, open(path, "w"), indent=0)
    print("wrote", path, len(cases), "cases;", {k: sum(1 for c in cases if k in c) for k in ("g_assign", "g_add", "H_assign", "H_add")})


if __name__ == "__main__":
    main()
