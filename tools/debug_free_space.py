"""diagnostic: product vs oracle free-space solve, per-level differences after the first call (GPU box)"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import octree_mg_b200 as om
from oracle import oracle as O
import problems as P
import test_free_space as T


import functools


def run(n, box, frac, fmg=True, tail_cells=None, calls=1, n_cpu=1):
    orc = T.build(O.OracleMg, n, box, **({"n_cpu": n_cpu} if n_cpu > 1 else {}))
    mg = T.build(om.Mg if n_cpu == 1 else functools.partial(om.MgGroup, n_cpu), n, box)
    if tail_cells is not None:
        mg.set_option("tail_cells", tail_cells)
    h_o = [orc.mg_poisson_free_3d(i == 0, frac, fmgcycle=fmg) for i in range(calls)]
    h_g = [mg.mg_poisson_free_3d(i == 0, frac, fmgcycle=fmg) for i in range(calls)]
    print("case ranks=%d n=%s box=%d frac=%s fmg=%s tail_cells=%s  fft_lvl %d/%d  res %r vs %r" % (n_cpu, n, box, frac, fmg, tail_cells, orc.free_space_fft_lvl, mg.free_space_fft_lvl, h_o, h_g))
    for lvl in range(orc.lowest_lvl, orc.highest_lvl + 1):
        d = {}
        for iv, name in ((1, "phi"), (2, "rhs")):
            dd = 0.0; mm = 0.0; dg = 0.0
            for id_ in orc.ids(lvl):
                a = orc.cc(id_)[iv - 1]; b = mg.cc(id_)[iv - 1]
                dd = max(dd, float(np.abs(a[1:-1, 1:-1, 1:-1] - b[1:-1, 1:-1, 1:-1]).max())); mm = max(mm, float(np.abs(a[1:-1, 1:-1, 1:-1]).max()))
                nc = a.shape[0] - 2
                s = slice(1, nc + 1)
                for sl in [(s, s, 0), (s, s, nc + 1), (s, 0, s), (s, nc + 1, s), (0, s, s), (nc + 1, s, s)]:
                    dg = max(dg, float(np.abs(a[sl] - b[sl]).max()))
            d[name] = (dd, dg, mm)
        print("   lvl %2d nc %2d boxes %3d  " % (lvl, orc.box_size_lvl(lvl), len(orc.ids(lvl))), "  ".join("%s: interior %.3e ghost %.3e (max %.3e)" % (k, *v) for k, v in d.items()))


run(32, 8, 1.0, n_cpu=2)
run(32, 16, 1.0, n_cpu=2)
run(32, 8, 0.15, n_cpu=2)
run(64, 16, 0.15, n_cpu=2)
run(32, 8, 0.15, n_cpu=4)
run(32, 8, 0.15, fmg=False, n_cpu=2)
