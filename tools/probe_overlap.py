#!/usr/bin/env python
"""Timing probe (torchrun, N GPUs): per level, colour half-sweeps alone, halo exchanges alone, and smooth_boxes with the
exchange serial (overlap=0) or overlapped with the interior boxes (overlap=1).  Plain stream launches (no graph)."""
import ctypes as C
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import torch.distributed as dist
import octree_mg_b200 as om
import problems as P
from bench import domain_for

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dims = domain_for(512, world)
mg = om.Mg(P.mg_laplacian, P.mg_smoother_gsrb, n_cpu=world, my_rank=rank, device=local, gpu_resident=True)
mg.set_int("host_cc", 0)
mg.set_int("coarse_collapse_boxes", 128)
for nb in range(1, 7):
    mg.set_bc(nb, P.mg_iphi, P.mg_bc_dirichlet, 0.0)
mg.mg_set_methods()
mg.mg_build_rectangle(dims, 16, [1.0 / dims[0]] * 3, [0.0] * 3, [False] * 3)
mg.mg_load_balance()
mg.mg_load_balance_parents()
mg.mg_allocate_storage()
uid = C.create_string_buffer(128)
if rank == 0:
    mg.G.omg_nccl_unique_id(uid)
t = torch.frombuffer(bytearray(uid.raw), dtype=torch.uint8).cuda()
dist.broadcast(t, 0)
uid = C.create_string_buffer(bytes(t.cpu().numpy().tobytes()), 128)
mg._gck(mg.G.omg_comm_init(mg.gpu, uid, rank, world))
G, gpu = mg.G, mg.gpu
mg._gck(G.omg_set_var_const(gpu, P.mg_irhs, -20, 20, 1.0, 1))
mg._gck(G.omg_set_var_const(gpu, P.mg_iphi, -20, 20, 0.0, 0))
mg.mg_fas_vcycle(want_res=False)
NC = 25


def run(what, lvl, overlap):
    mg.set_option("overlap", overlap)
    out = C.c_double(0.0)
    mg._gck(G.omg_debug_step(gpu, what.encode(), lvl, 2, C.byref(out)))      # warm-up
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    mg._gck(G.omg_debug_step(gpu, what.encode(), lvl, NC, C.byref(out)))
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], device="cuda")
    dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    return float(dt.item()) / (2 * NC) * 1e6      # us per half-sweep


for lvl in (1, 0, -1):
    a = run("smooth_only", lvl, 0)
    b = run("halo_only", lvl, 0)
    c = run("smooth_boxes", lvl, 0)
    d = run("smooth_boxes", lvl, 1)
    if rank == 0:
        print("lvl %2d: half-sweep %.1f us, halo exchange alone %.1f us, sweep+halo serial %.1f us, overlapped %.1f us" % (lvl, a, b, c, d), flush=True)
