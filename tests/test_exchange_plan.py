"""CPU tests of the multi-rank host logic: the static exchange plans (omg_plan.cpp through omg_exchange_plan)
that replace the reference's run-time pack + sort (src/m_communication.f90:68-88).  No device is needed:
the tree and the Morton partition come from the host mirror, the plan from the C ABI.

 * in one process: for every pair of ranks, what r sends to p is exactly what p expects from r, in the same order;
   the counts agree with the oracle's own buffer sizing (the reference's comm_*%n_send / n_recv dry run);
 * with torch.distributed (gloo, world_size 2): each rank computes only its own plan, the ranks exchange their
   key lists and a payload routed by those keys arrives where the receiver expects it."""
import os
import subprocess
import sys

import numpy as np
import pytest

import problems as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build_tree(product, n_cpu, rank, amr=False, lb="default"):
    mg = product.Mg(P.mg_laplacian, P.mg_smoother_gsrb, n_cpu=n_cpu, my_rank=rank)
    mg.mg_set_methods()
    if amr:
        # same as P.build_amr_tree but without allocating device storage
        alloc = mg.mg_allocate_storage
        mg.mg_allocate_storage = lambda: None
        P.build_amr_tree(mg, 3, [32] * 3, 8, [1.0 / 32] * 3)
        mg.mg_allocate_storage = alloc
    else:
        mg.mg_build_rectangle([64, 32, 32], 8, [1.0 / 64] * 3, [0.0] * 3, [False, True, False])
        if lb == "simple":
            mg.mg_load_balance_simple()
        else:
            mg.mg_load_balance()
            if lb == "parents":
                mg.mg_load_balance_parents()
    return mg


@pytest.mark.parametrize("n_cpu,amr,lb", [(2, False, "default"), (3, False, "simple"), (8, False, "parents"), (4, True, "default")])
def test_plans_are_symmetric(product, n_cpu, amr, lb):
    mgs = [build_tree(product, n_cpu, r, amr, lb) for r in range(n_cpu)]
    total = 0
    for phase in (0, 1, 2):
        for lvl in range(mgs[0].lowest_lvl, mgs[0].highest_lvl + 1):
            for r in range(n_cpu):
                for p in range(n_cpu):
                    if p == r:
                        continue
                    send, _ = mgs[r].exchange_plan(phase, lvl, p)
                    _, recv = mgs[p].exchange_plan(phase, lvl, r)
                    assert send == recv, (phase, lvl, r, p)
                    total += len(send)
    assert total > 0


def test_ghost_plan_matches_tree(product):
    """every face whose neighbour lives on another rank appears exactly once, keyed 6*id + face"""
    n_cpu = 4
    mg = build_tree(product, n_cpu, 1)
    t = mg.tree_export()
    for lvl in range(mg.lowest_lvl, mg.highest_lvl + 1):
        expect = {}
        for id_ in mg.ids(lvl):
            if t[id_ - 1, 0] != 1:
                continue
            for f in range(6):
                nb = t[id_ - 1, 14 + f]
                if nb > 0 and t[nb - 1, 0] != 1:
                    expect.setdefault(int(t[nb - 1, 0]), []).append(2 * (6 * id_ + f))
        for p in range(n_cpu):
            if p == 1:
                continue
            _, recv = mg.exchange_plan(0, lvl, p)
            assert recv == sorted(expect.get(p, [])), (lvl, p)


GLOO_WORKER = r"""
import os, sys
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import numpy as np
import torch, torch.distributed as dist
import octree_mg_b200 as om
from test_exchange_plan import build_tree
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=world)
mg = build_tree(om, world, rank, amr=%(amr)r)
peer = 1 - rank
checked = 0
for phase in (0, 1, 2):
    for lvl in range(mg.lowest_lvl, mg.highest_lvl + 1):
        send, recv = mg.exchange_plan(phase, lvl, peer)
        # counts first (what the reference precomputes in comm_*%%n_recv), then the routed payload
        n = torch.tensor([len(send)]); m = torch.zeros(1, dtype=torch.long)
        reqs = [dist.isend(n, peer), dist.irecv(m, peer)]
        [q.wait() for q in reqs]
        assert int(m) == len(recv), (phase, lvl, int(m), len(recv))
        if len(send) + len(recv) == 0:
            continue
        out = torch.tensor(send, dtype=torch.long) * 3 + 1 if send else torch.zeros(0, dtype=torch.long)
        inp = torch.zeros(len(recv), dtype=torch.long)
        reqs = ([dist.isend(out, peer)] if send else []) + ([dist.irecv(inp, peer)] if recv else [])
        [q.wait() for q in reqs]
        assert inp.tolist() == [k * 3 + 1 for k in recv], (phase, lvl)
        checked += len(recv)
tot = torch.tensor([checked]); dist.all_reduce(tot)
assert int(tot) > 0
dist.barrier()
sys.stdout.write("rank%%d-ok\n" %% rank); sys.stdout.flush()
"""


@pytest.mark.parametrize("amr", [False, True])
def test_gloo_two_ranks(product, amr, tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(GLOO_WORKER % {"root": ROOT, "amr": amr})
    port = 29500 + (os.getpid() + (7 if amr else 0)) % 2000
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), str(script)]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:]
    assert "rank0-ok" in r.stdout and "rank1-ok" in r.stdout, r.stdout[-2000:]
