"""A handful of bitwise product-vs-oracle checks that finish in seconds on the GPU box (no torch import):
the multi-box tail fast path (non-cubic grids, constant boundary conditions) and the single-box one."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import functools

import octree_mg_b200 as om
import problems as P
from compare import assert_state_equal
from oracle import oracle as O

t0 = time.time()
ok = True
cases = [("lpl 64x32x32 box16 dirichlet0", dict(n=(64, 32, 32), box=16, bc="dirichlet0"), {}),
         ("lpl 32x64x32 box8 mixed", dict(n=(32, 64, 32), box=8, bc="mixed"), {}),
         ("helmh 64x32x32 box16 neumann0", dict(n=(64, 32, 32), box=16, bc="neumann0", operator_type=P.mg_helmholtz, lam=1.5), {}),
         ("lpl 32^3 box16 continuous_x", dict(n=32, box=16, bc="continuous_x"), {}),
         ("lpl 64x32x32 box16 dirichlet0, 2 ranks", dict(n=(64, 32, 32), box=16, bc="dirichlet0"), dict(n_cpu=2))]
for name, kw, mr in cases:
    try:
        n, box = kw.pop("n"), kw.pop("box")
        orc, _ = P.uniform_problem(O.OracleMg, n, box, **kw, **mr)
        cls = functools.partial(om.MgGroup, mr["n_cpu"]) if mr else om.Mg
        mg, _ = P.uniform_problem(cls, n, box, **kw)
        ho = [orc.mg_fas_fmg(False)] + [orc.mg_fas_vcycle() for _ in range(3)]
        hg = [mg.mg_fas_fmg(False)] + [mg.mg_fas_vcycle() for _ in range(3)]
        assert ho == hg, (ho, hg)
        assert_state_equal(orc, mg)
        print("OK  ", name, "%.1fs" % (time.time() - t0), flush=True)
    except BaseException as e:  # noqa: BLE001
        ok = False
        print("FAIL", name, repr(e)[:300], flush=True)
print("quick_check", "PASSED" if ok else "FAILED", "%.1fs" % (time.time() - t0))
