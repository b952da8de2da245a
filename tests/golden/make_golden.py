"""Generate tests/golden/golden.json from the CPU oracle.

The reference ships no golden vectors and cannot be compiled in this image (F2), so
these fixtures are ORACLE output, committed to (a) detect regressions of the oracle and (b) let the GPU
tests compare against numbers that do not depend on recomputing the oracle at test time.  Residuals
are stored as hex floats (exact); phi through a position-weighted checksum plus a sampled slice.

Regenerate with:  python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import problems as P  # noqa: E402

CASES = {
    "uniform32_box8_vcycle": dict(kind="uniform", n=32, box=8, fmg=False, its=4),
    "uniform32_box8_fmg": dict(kind="uniform", n=32, box=8, fmg=True, its=4),
    "uniform64_box16_fmg": dict(kind="uniform", n=64, box=16, fmg=True, its=3),
    "perf64_box16_vcycle": dict(kind="perf", n=64, box=16, fmg=False, its=4),
    "helmholtz32_neumann_fmg": dict(kind="uniform", n=32, box=8, fmg=True, its=3, op=P.mg_helmholtz, bc="neumann0", lam=1.0),
    "vlaplacian32_mixed_fmg": dict(kind="uniform", n=32, box=8, fmg=True, its=3, op=P.mg_vlaplacian, bc="mixed"),
    "refinement3_box8_fmg": dict(kind="amr", n=32, box=8, levels=3, fmg=True, its=3),
    "uniform32_box8_gs_vcycle": dict(kind="uniform", n=32, box=8, fmg=False, its=4, smoother=P.mg_smoother_gs),
    "ahelmholtz32_mixed_fmg": dict(kind="uniform", n=32, box=8, fmg=True, its=3, op=P.mg_ahelmholtz, bc="mixed", lam=0.5),
    # BASELINE.json configs[0]: test_uniform_grid 128^3, box 16, 10 iterations, FMG and V-cycle, GSRB and the program's own GS
    "config0_uniform128_box16_fmg": dict(kind="uniform", n=128, box=16, fmg=True, its=10),
    "config0_uniform128_box16_vcycle": dict(kind="uniform", n=128, box=16, fmg=False, its=10),
    "config0_uniform128_box16_gs_fmg": dict(kind="uniform", n=128, box=16, fmg=True, its=10, smoother=P.mg_smoother_gs),
    "noncubic64x32x32_box16_fmg": dict(kind="uniform", n=[64, 32, 32], box=16, fmg=True, its=3, bc="dirichlet0"),
    # implicit diffusion steps (src/m_diffusion.f90): residuals = [last max residual, number of V-cycles added] per step
    "diffusion32_box8_order1": dict(kind="diffusion", n=32, box=8, order=1, steps=2, op=P.mg_helmholtz),
    "diffusion32_box8_order2": dict(kind="diffusion", n=32, box=8, order=2, steps=2, op=P.mg_helmholtz),
    "diffusion32_box8_vcoeff_order2": dict(kind="diffusion", n=32, box=8, order=2, steps=1, op=P.mg_vhelmholtz),
}


def checksum(mg):
    """exact, order-fixed summary of phi on all leaves: hex of sum(w * phi) per box folded in id order,
    where w is a fixed position weight; plus min / max"""
    acc = 0.0
    lo, hi = np.inf, -np.inf
    for lvl in range(mg.lowest_lvl, mg.highest_lvl + 1):
        for id_ in mg.leaves(lvl):
            a = mg.cc(id_)[0, 1:-1, 1:-1, 1:-1]
            w = 1.0 + (np.arange(a.size).reshape(a.shape) % 7) * 0.125
            acc = acc * 0.5 + float(np.sum((a * w).ravel()[::1].cumsum()[-1]))
            lo, hi = min(lo, float(a.min())), max(hi, float(a.max()))
    return [float(acc).hex(), float(lo).hex(), float(hi).hex()]


def build(cls, name, **kw):
    c = CASES[name]
    if c["kind"] == "uniform":
        mg, _ = P.uniform_problem(cls, c["n"], c["box"], operator_type=c.get("op", P.mg_laplacian),
                                  smoother=c.get("smoother", P.mg_smoother_gsrb), bc=c.get("bc", "callback"),
                                  lam=c.get("lam", 0.0), **kw)
    elif c["kind"] == "perf":
        mg = P.performance_problem(cls, c["n"], c["box"], **kw)
    elif c["kind"] == "diffusion":
        mg = P.diffusion_problem(cls, c["n"], c["box"], operator_type=c["op"], **kw)
    else:
        mg, _ = P.refinement_problem(cls, c["levels"], c["box"], c["n"], **kw)
    return mg


def run_case(cls, name, **kw):
    c = CASES[name]
    mg = build(cls, name, **kw)
    if c["kind"] == "diffusion":
        res = []
        for _ in range(c["steps"]):
            r, n = (mg.diffusion_solve(2.0e-3, 1.5, c["order"], 1e-9) if c["op"] == P.mg_helmholtz
                    else mg.diffusion_solve_vcoeff(2.0e-3, c["order"], 1e-9))
            res += [r, float(n)]
    else:
        res = P.run_cycles(mg, c["its"], c["fmg"])
    return dict(name=name, residuals=[float(r).hex() for r in res], phi_checksum=checksum(mg))


if __name__ == "__main__":
    from oracle import oracle as O
    out = dict(generator="tests/golden/make_golden.py (oracle/omg_oracle.cpp, g++ -O2 -ffp-contract=off)",
               cases=[run_case(O.OracleMg, name) for name in CASES])
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", len(out["cases"]), "cases")
