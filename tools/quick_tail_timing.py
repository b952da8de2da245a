"""V-cycle time on a non-cubic grid (two boxes on every coarse level) with and without the tail fast path."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import octree_mg_b200 as om
import problems as P

for dims in ((256, 128, 128), (128, 128, 128)):
    mg = P.performance_problem(om.Mg, dims, 16, gpu_resident=True)
    mg.mg_gpu_push(P.mg_irhs)
    mg.mg_gpu_push(P.mg_iphi)
    for opt in (1, 0, 1):
        mg.set_option("tail_simple", opt)
        mg.mg_fas_vcycle(want_res=False)
        mg.fas_vcycles(3, want_res=False)
        _, ms = mg.fas_vcycles(20, want_res=False)
        print(dims, "tail_simple", opt, "%.4f ms per V-cycle" % (ms / 20), flush=True)
