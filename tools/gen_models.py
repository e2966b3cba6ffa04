#!/usr/bin/env python
"""Generates the straight-line model code of the two quadcopter problems with sympy
(the reference generates its problem functions with CasADi: problems/generate_problem.py,
alpaqa/python/alpaqa/casadi_generator/__init__.py:50-190; CasADi is not available here).

Math spec (reference, paths relative to /root/reference/python/alpaqa_mpc_benchmarks):
  problems/quadcopter.py:17-47            f_c of the simplified quadcopter, ZYX rotation matrix
  problems/realistic_quadcopter.py:17-52  f_c of the realistic quadcopter, ARB and Omega^-1 omega
  problems/quadcopter.py:62-75            stage/terminal constraint c(x) (same in both files)

For f_c: R^9 x R^4 -> R^9 and c: R^9 -> R^4 it emits, as `PANTR_HD inline` functions,
  <f>(z) ; <f>_jvp(z, dz) = J dz ; <f>_vjp(z, w) = J^T w ; <f>_hvjp(z, w, dz) = d/dz[J^T w] dz
with the sines/cosines of the three angles passed in as `sc[6]` (they are computed once per stage
point by pantr_sincos and cached by the rollout).  The same header is compiled into the CUDA
library and into the CPU oracle, so both evaluate identical operation sequences.  Sums of products
are emitted as explicit fma() chains (the compilers' own contraction is off on both sides), which
halves the FP64 instruction count of the derivative code.  Besides the functions the header
carries FC_CONST_COLS (the columns of d f_c / d(x, u) that are constant AND only feed constant
columns: their columns of the RK4 step Jacobian are the same for every step and every instance)
and generated/flop_counts.json lists the exact operation count of every emitted function
(add = mul = div = 1, fma = 2), the numerator of the roofline.

Usage: python tools/gen_models.py   (writes pantr_cdc2023_experiments_b200/csrc/generated/*.h)
"""
import os

import sympy as sp
from sympy.printing.c import C99CodePrinter

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "pantr_cdc2023_experiments_b200", "csrc", "generated")

x = sp.symbols("x0:9", real=True)
u = sp.symbols("u0:4", real=True)
dx = sp.symbols("dx0:9", real=True)
du = sp.symbols("du0:4", real=True)
w = sp.symbols("w0:9", real=True)
yh = sp.symbols("yh0:4", real=True)
S = sp.symbols("sc0 sc2 sc4", real=True)   # sines of x6, x7, x8
Cc = sp.symbols("sc1 sc3 sc5", real=True)  # cosines
TRIG = {}
for i in range(3):
    TRIG[sp.sin(x[6 + i])] = S[i]
    TRIG[sp.cos(x[6 + i])] = Cc[i]
G = sp.Float(9.81)


def fc_quadcopter():
    """quadcopter.py:30-47"""
    p, v, th = x[0:3], x[3:6], x[6:9]
    at, om = u[0], u[1:4]
    cr, sr, cp, sp_, cy, sy = sp.cos(th[0]), sp.sin(th[0]), sp.cos(th[1]), sp.sin(th[1]), sp.cos(th[2]), sp.sin(th[2])
    R = sp.Matrix([[cy * cp, cy * sp_ * sr - sy * cr, cy * sp_ * cr + sy * sr],
                   [sy * cp, sy * sp_ * sr + cy * cr, sy * sp_ * cr - cy * sr],
                   [-sp_, cp * sr, cp * cr]])
    a = R * sp.Matrix([0, 0, at]) + sp.Matrix([0, 0, -G])
    return list(v) + list(a) + list(om)


def fc_realistic_quadcopter():
    """realistic_quadcopter.py:29-52 (orient = (phi, theta, psi) = x6, x7, x8)"""
    v = x[3:6]
    phi, th, psi = x[6], x[7], x[8]
    at, om = u[0], sp.Matrix(u[1:4])
    cpsi, cphi, cth = sp.cos(psi), sp.cos(phi), sp.cos(th)
    spsi, sphi, sth = sp.sin(psi), sp.sin(phi), sp.sin(th)
    ARB = sp.Matrix([[cpsi * cth - sphi * spsi * sth, -cphi * spsi, cpsi * sth + cth * sphi * spsi],
                     [cth * spsi + cpsi * sphi * sth, cphi * cpsi, spsi * sth - cpsi * cth * sphi],
                     [-cphi * sth, sphi, cphi * cth]])
    Om = sp.Matrix([[cth, 0, -cphi * sth], [0, 1, sphi], [sth, 0, cphi * cth]])
    d_orient = sp.simplify(Om.LUsolve(om))
    # keep only sin/cos of the angles (they are passed in as sc[]): no tan in the emitted code
    d_orient = d_orient.applyfunc(lambda e: e.replace(sp.tan, lambda a: sp.sin(a) / sp.cos(a)))
    a = ARB * sp.Matrix([0, 0, at]) + sp.Matrix([0, 0, -G])
    return list(v) + list(a) + list(d_orient)


def constraints():
    """quadcopter.py:62-69"""
    return [x[6], x[7], sp.cos(x[6]) * sp.cos(x[7]), sp.Float(0.1) ** 2 - x[0] ** 2 - x[1] ** 2]


class Printer(C99CodePrinter):
    """C printer that writes a sum with product terms as a chain of fma() and counts the operations it emits."""
    flops = 0

    def _print_Pow(self, e):
        if e.exp.is_Integer and 1 < int(e.exp) <= 4:
            b = self._print(e.base)
            if not e.base.is_Atom:
                b = "(" + b + ")"
            Printer.flops += int(e.exp) - 1
            return "(" + "*".join([b] * int(e.exp)) + ")"
        if e.exp == -1:
            Printer.flops += 1
            return "(1.0/" + self.parenthesize(e.base, 100) + ")"
        if e.exp.is_Integer and -4 <= int(e.exp) < -1:
            b = self._print(e.base)
            if not e.base.is_Atom:
                b = "(" + b + ")"
            Printer.flops += -int(e.exp)
            return "(1.0/(" + "*".join([b] * (-int(e.exp))) + "))"
        return super()._print_Pow(e)

    def _print_Mul(self, e):
        c, rest = e.as_coeff_Mul()
        nf = len(sp.Mul.make_args(rest)) + (0 if c in (1, -1) else 1)
        Printer.flops += max(nf - 1, 0)
        return super()._print_Mul(e)

    def _print_Integer(self, e):
        return f"{int(e)}.0" if int(e) >= 0 else f"({int(e)}.0)"

    def _split(self, t):
        """product term -> (a, b) with t = a * b, both printed"""
        c, rest = t.as_coeff_Mul()
        fac = list(sp.Mul.make_args(rest)) if rest != 1 else []
        if c not in (1, -1):
            fac = [c] + fac
            sign = ""
        else:
            sign = "-" if c == -1 else ""
        a, b = fac[0], sp.Mul(*fac[1:])
        return sign + self.parenthesize(a, 100), self._print(b)

    @staticmethod
    def _is_prod(t):
        c, rest = t.as_coeff_Mul()
        nf = (len(sp.Mul.make_args(rest)) if rest != 1 else 0) + (0 if c in (1, -1) else 1)
        return nf >= 2

    def _print_Add(self, e):
        terms = list(e.as_ordered_terms())
        prods = [t for t in terms if self._is_prod(t)]
        plain = [t for t in terms if not self._is_prod(t)]
        if not prods:
            Printer.flops += len(terms) - 1
            return super()._print_Add(e)
        if plain:
            acc = self._print(plain[0]) if len(plain) == 1 else super()._print_Add(sp.Add(*plain, evaluate=False))
            Printer.flops += len(plain) - 1
            rest = prods
        else:
            acc = self._print(prods[0])
            rest = prods[1:]
        for t in rest:
            a, b = self._split(t)
            acc = f"fma({a}, {b}, {acc})"
            Printer.flops += 2
        return acc


PR = Printer({"precision": 17})


FLOPS = {}  # emitted function name -> exact operation count (add = mul = div = 1, fma = 2)
ALL_FLOPS = {}


def emit(name, args, outs, exprs):
    """args: list of (c_name, symbols); outs: list of (c_name, n); exprs: flat list of sympy exprs."""
    Printer.flops = 0
    exprs = [sp.sympify(e).subs(TRIG) for e in exprs]
    for e in exprs:
        assert not e.has(sp.sin) and not e.has(sp.cos) and not e.has(sp.tan), e
    repl, red = sp.cse(exprs, symbols=sp.numbered_symbols("t"), order="canonical")
    used = set().union(*[e.free_symbols for e in exprs]) if exprs else set()
    sig = ", ".join(f"const double *{a}" for a, _ in args) + ", " + ", ".join(f"double *{o}" for o, _ in outs)
    lines = [f"PANTR_HD void {name}({sig}) {{"]
    for a, syms in args:
        for i, s in enumerate(syms):
            if s in used:
                lines.append(f"    const double {s} = {a}[{i}];")
    for a, _ in args:
        lines.append(f"    (void){a};")
    for s, e in repl:
        lines.append(f"    const double {s} = {PR.doprint(e)};")
    k = 0
    for o, n in outs:
        for i in range(n):
            lines.append(f"    {o}[{i}] = {PR.doprint(red[k])};")
            k += 1
    lines.append("}")
    FLOPS[name] = Printer.flops
    lines.insert(0, f"// {Printer.flops} flops")
    return "\n".join(lines)


def family(prefix, F, zvars, dzvars, zargs, dzargs, wsyms, wname, outs):
    """value, jvp, vjp, hvjp of F(z).  outs: list of (name, n) splitting the z-gradient outputs."""
    F = [sp.sympify(f) for f in F]
    n = len(F)
    J = sp.Matrix(F).jacobian(sp.Matrix(zvars))
    sc_arg = [("sc", list(S[:1]) + list(Cc[:1]) + list(S[1:2]) + list(Cc[1:2]) + list(S[2:]) + list(Cc[2:]))]
    parts = []
    parts.append(emit(prefix, zargs + sc_arg, [("k", n)], F))
    parts.append(emit(prefix + "_jvp", zargs + sc_arg + dzargs, [("dk", n)], list(J * sp.Matrix(dzvars))))
    g = sum(wi * fi for wi, fi in zip(wsyms, F))
    grad = [sp.diff(g, z) for z in zvars]
    parts.append(emit(prefix + "_vjp", zargs + sc_arg + [(wname, wsyms)], [("a" + o, k) for o, k in outs], grad))
    hv = [sum(sp.diff(gi, zj) * dzj for zj, dzj in zip(zvars, dzvars)) for gi in grad]
    parts.append(emit(prefix + "_hvjp", zargs + sc_arg + [(wname, wsyms)] + dzargs, [("b" + o, k) for o, k in outs], hv))
    return "\n\n".join(parts)


def const_cols(fc):
    """Columns c of J = d f_c / d(x, u) that are constant and whose nonzero rows are themselves constant columns:
    a unit tangent in such a column stays a constant vector through every RK4 stage."""
    z = list(x) + list(u)
    J = sp.Matrix(fc).jacobian(sp.Matrix(z))
    cst = {c for c in range(len(z)) if all(J[r, c].is_number for r in range(J.rows))}
    while True:
        bad = {c for c in cst if any(J[r, c] != 0 and r not in cst for r in range(J.rows))}
        if not bad:
            break
        cst -= bad
    return sum(1 << c for c in cst)


def write_model(ns, fc, ref):
    FLOPS.clear()
    body = family("fc", fc, list(x) + list(u), list(dx) + list(du), [("x", x), ("u", u)], [("dx", dx), ("du", du)],
                  w, "w", [("x", 9), ("u", 4)])
    body += "\n\n" + family("con", constraints(), list(x), list(dx), [("x", x)], [("dx", dx)], yh, "yh", [("x", 9)])
    text = (f"// GENERATED by tools/gen_models.py (sympy {sp.__version__}) -- do not edit.\n"
            f"// Model: {ref}\n"
            "// fc: continuous dynamics f_c(x, u); con: stage/terminal constraint c(x).  sc = {sin x6, cos x6,\n"
            "// sin x7, cos x7, sin x8, cos x8}.  *_jvp = J dz, *_vjp = J^T w, *_hvjp = d/dz[J^T w] dz.\n"
            "#pragma once\n#include \"../pantr_hd.h\"\n\n"
            f"namespace pantr {{\nnamespace {ns} {{\n\n"
            "// columns of d f_c / d(x, u) (bit c = column c of z = (x, u)) along which a unit tangent stays constant\n"
            f"constexpr unsigned FC_CONST_COLS = {hex(const_cols(fc))};\n\n"
            f"{body}\n\n}} // namespace {ns}\n}} // namespace pantr\n")
    ALL_FLOPS[ns] = dict(FLOPS)
    path = os.path.join(OUT, f"{ns}.h")
    with open(path, "w") as fh:
        fh.write(text)
    print("wrote", path, len(text.splitlines()), "lines")


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    write_model("quadcopter_gen", fc_quadcopter(), "python/alpaqa_mpc_benchmarks/problems/quadcopter.py:17-75")
    write_model("realistic_quadcopter_gen", fc_realistic_quadcopter(),
                "python/alpaqa_mpc_benchmarks/problems/realistic_quadcopter.py:17-75")
    import json
    with open(os.path.join(OUT, "flop_counts.json"), "w") as fh:
        json.dump(ALL_FLOPS, fh, indent=1, sort_keys=True)
        fh.write("\n")
