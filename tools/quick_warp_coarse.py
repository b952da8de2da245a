"""A/B of option warp_coarse (coarse solve of the single 2^3 box by one warp in registers) on problems whose cycle the tail
dominates: V-cycle time, coarse iteration count and bitwise equality of the residual histories of the two settings."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import octree_mg_b200 as om
import problems as P

D = (256, 128, 128)      # non-cubic: the coarsest level has two 2^3 boxes
CASES = [("laplacian dirichlet0 rhs=1 256x128x128", 0, lambda: P.performance_problem(om.Mg, D, 16, gpu_resident=True)),
         ("helmholtz neumann0 256x128x128", 0, lambda: P.uniform_problem(om.Mg, D, 16, operator_type=P.mg_helmholtz, bc="neumann0", lam=1.0, gpu_resident=True)[0]),
         ("laplacian dirichlet0 rhs=1 128^3", 0, lambda: P.performance_problem(om.Mg, 128, 16, gpu_resident=True)),
         ("helmholtz neumann0 128^3", 0, lambda: P.uniform_problem(om.Mg, 128, 16, operator_type=P.mg_helmholtz, bc="neumann0", lam=1.0, gpu_resident=True)[0]),
         ("vlaplacian mixed 128^3", 1, lambda: P.uniform_problem(om.Mg, 128, 16, operator_type=P.mg_vlaplacian, bc="mixed", gpu_resident=True)[0])]
for name, ne, make in CASES:
    hist = {}
    for opt in (1, 0):
        mg = make()
        mg.set_option("warp_coarse", opt)
        for iv in [P.mg_iphi, P.mg_irhs] + [P.mg_iveps + d for d in range(ne)]:
            mg.mg_gpu_push(iv)
        h = [mg.mg_fas_vcycle() for _ in range(3)]
        mg.fas_vcycles(3, want_res=False)
        _, ms = mg.fas_vcycles(20, want_res=False)
        hist[opt] = [x.hex() for x in h]
        print(name, "warp_coarse", opt, "%.4f ms per V-cycle" % (ms / 20), "coarse iterations of the last cycle", mg.stat("coarse_iters"), flush=True)
        mg.mg_deallocate_storage()
    print(name, "residual histories bitwise equal:", hist[0] == hist[1], hist[1], flush=True)
