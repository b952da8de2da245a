#!/usr/bin/env python
"""bench.py -- 3D Poisson FAS V-cycle throughput (GDOF/s) on B200, with the roofline of the dominant kernel,
an end-to-end number through the C ABI with host buffers, a strong-scaling record of the 1024^3 domain, a bitwise
parity check of the data plane actually timed, and the CPU oracle timed beside it.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--cells 512] [--box 16]

A step is one standalone V(2,2) GSRB cycle (mg_fas_vcycle, src/m_multigrid.f90:150-243) of the
tests/test_performance.f90 problem (rhs = 1, phi0 = 0, homogeneous Dirichlet), called as in that program's timed
loop (:68-76: `call mg_fas_vcycle(mg)`, no max_res; --max-res adds the all-level residual pass).
N = 1: BASELINE.json configs[1] (512^3, box 16).  N > 1: weak scaling, 512^3 per GPU along the reference's
Morton partition (N = 8 is configs[4], 1024^3).  Every line also carries
  "strong":       the SAME 1024^3 domain (configs[4]) timed on the N GPUs of this run (N = 1 included),
  "parity_check": FMG + 3 V-cycles of a 256^3 test_uniform_grid problem on the communicator of this run (all N
                  processes: CUDA-IPC puts + NCCL), residual history and a checksum of phi compared bit for bit
                  with the CPU oracle on rank 0; a mismatch makes the run exit non-zero.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

BYTES_PER_DOF_VCYCLE = 155.4            # Laplacian V(2,2); + 18.3 when max_res is requested
BYTES_PER_DOF_MAXRES = 18.3
SMOOTH_BYTES_PER_CELL = 12.0            # one colour half-sweep launched on its own
CPU_NS_PER_CELL_CYCLE = 80.0            # oracle, one thread: pessimistic estimate used only to bound the CPU arm's run time


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(n, box):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE finest-level smoother launch, from the newest committed
    `ncu --set full` capture of this very command (profiles/, tools/ncu_summary.py); None for other workloads."""
    if (n, box) != (512, 16):
        return None, None
    cands = sorted(f for f in os.listdir(os.path.join(ROOT, "profiles")) if f.endswith("ncu_smooth_finest.txt"))
    if not cands:
        return None, None
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    tot = 0.0
    for line in open(os.path.join(ROOT, "profiles", cands[-1])):
        f = line.split()
        if len(f) >= 3 and f[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            tot += float(f[1]) * unit.get(f[2], 1.0)
    return (tot or None), "profiles/" + cands[-1]


class ClockSampler:
    def __init__(self, index=0):
        self.rows, self.index, self.proc = [], index, None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index),
                 "--query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
                 "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
                 "clocks_event_reasons.sw_power_cap", "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        loaded = [r for r in self.rows if len(r) > 2 and r[0].isdigit()]
        try:   # under load = power draw in the upper half of the observed range
            pw = [float(r[2]) for r in loaded]
            cut = 0.5 * (min(pw) + max(pw))
            loaded = [r for r, w in zip(loaded, pw) if w >= cut] or loaded
        except Exception:
            pass
        sm = sorted(int(r[0]) for r in loaded)
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def domain_for(n, world):
    """weak scaling: n^3 per GPU; the Morton-contiguous partition of mg_load_balance gives each rank one cube"""
    dims = [n, n, n]
    w, d = world, 0
    while w > 1:
        dims[d % 3] *= 2
        w //= 2
        d += 1
    return dims


def workload_text(dims, box, max_res):
    return ("test_performance %s box %d: 7-point Laplacian, rhs=1, Dirichlet 0, standalone V(2,2) GSRB cycles%s"
            % ("x".join(map(str, dims)), box, " with max_res" if max_res else " (mg_fas_vcycle(mg) as in test_performance.f90:68-76)"))


def host_threads():
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    ranks = 1
    while ranks * 2 <= cores:
        ranks *= 2
    return ranks


def mem_available_bytes():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                return int(line.split()[1]) * 1024
    except Exception:
        pass
    return 16 << 30


def cpu_reference_line(dims, box, steps, warmup, max_res, budget_s):
    """The CPU arm: the oracle (C++ restatement of the reference; the Fortran/MPI reference cannot be built in this
    image) with one OpenMP thread per emulated MPI rank, all host cores.  It runs the workload it is given -- same
    domain, box size, cycle, steps and warm-up as the GPU arm -- unless the estimated run time exceeds budget_s or the
    domain does not fit the host's memory; then the cycle count / the domain are cut down and the record says exactly
    what ran ("ran": domain, steps, warmup)."""
    import problems as P
    from oracle import oracle as O
    ranks = host_threads()
    dims = list(dims)
    note = []
    # memory: 4 variables with ghost shells per box, all levels (x 8/7), + slack
    def footprint(d):
        return d[0] * d[1] * d[2] / box ** 3 * (8.0 / 7.0) * 4 * (box + 2) ** 3 * 8 * 1.15
    while footprint(dims) > 0.6 * mem_available_bytes() and max(dims) > 128:
        dims[dims.index(max(dims))] //= 2
        note.append("domain halved to fit host memory")
    def est(d, cycles):
        return d[0] * d[1] * d[2] * CPU_NS_PER_CELL_CYCLE * 1e-9 / ranks * cycles
    while est(dims, 4) > budget_s and max(dims) > 128:
        dims[dims.index(max(dims))] //= 2
        note.append("domain halved to bound the run time")
    w = max(0, warmup)
    k = max(1, steps)
    while est(dims, w + k) > budget_s and (w > 1 or k > 3):
        if w > 1:
            w -= 1
        else:
            k = max(3, k - 1)
    O.lib().orc_set_omp_threads(ranks)
    mg = P.performance_problem(O.OracleMg, dims, box, n_cpu=ranks)
    dof = dims[0] * dims[1] * dims[2]
    for _ in range(w):
        mg.mg_fas_vcycle(want_res=max_res)
    t0 = time.perf_counter()
    for _ in range(k):
        mg.mg_fas_vcycle(want_res=max_res)
    dt = time.perf_counter() - t0
    val = dof * k / dt / 1e9
    ran = {"domain": dims, "steps": k, "warmup": w}
    return {"value": val, "unit": "GDOF/s", "cores": ranks, "kind": "port",
            "sample": "%s box %d, %d timed V(2,2) GSRB cycles after %d warm-up%s, oracle/omg_oracle.cpp, %d OpenMP threads = emulated MPI ranks%s"
                      % ("x".join(map(str, dims)), box, k, w, " with max_res" if max_res else "", ranks,
                         ("; " + ", ".join(sorted(set(note)))) if note else ""),
            "ms_per_step": dt / k * 1e3, "ran": ran,
            "ref_style_unknowns_per_us": mg.n_boxes * box ** 3 * k / (dt * 1e6)}


def phi_checksum(mg, P):
    """order-independent exact checksum (mod 2^64) of phi on the owned leaves: bit patterns weighted by position and box id"""
    import numpy as np
    tot = np.uint64(0)
    with np.errstate(over="ignore"):
        for lvl in range(mg.lowest_lvl, mg.highest_lvl + 1):
            nc = mg.box_size_lvl(lvl)
            w = (np.arange(nc ** 3, dtype=np.uint64) * np.uint64(2) + np.uint64(1)) * np.uint64(0x9E3779B97F4A7C15)
            mine = mg.my_leaves(lvl) if hasattr(mg, "my_leaves") else mg.leaves(lvl)
            for id_ in mine:
                v = np.ascontiguousarray(mg.cc(id_, nc)[P.mg_iphi - 1, 1:-1, 1:-1, 1:-1]).view(np.uint64).ravel()
                v = np.where(v == np.uint64(1 << 63), np.uint64(0), v)      # -0.0 == +0.0
                tot = tot + (v * w).sum(dtype=np.uint64) * np.uint64(2 * id_ + 1)
    return int(tot)


def comm_init(mg, dist, rank, world):
    """mg_comm_init of the GPU library: NCCL unique id from rank 0, broadcast through torch.distributed"""
    import torch
    uid = C.create_string_buffer(128)
    if rank == 0:
        mg.G.omg_nccl_unique_id(uid)
    t = torch.frombuffer(bytearray(uid.raw), dtype=torch.uint8).cuda()
    dist.broadcast(t, 0)
    uid = C.create_string_buffer(bytes(t.cpu().numpy().tobytes()), 128)
    mg._gck(mg.G.omg_comm_init(mg.gpu, uid, rank, world))


def parity_check(om, P, n, box, rank, world, local, dist, collapse):
    """FMG + 3 V-cycles of the test_uniform_grid problem (n^3 in total, Dirichlet planes from the manufactured solution)
    on the communicator this run timed, against the oracle (rank 0): residual history and phi checksum, bit for bit"""
    import torch

    def setup(mg):
        if world > 1:
            mg.set_int("coarse_collapse_boxes", collapse)
    mg, _ = P.uniform_problem(om.Mg, n, box, load_balance="parents", setup=setup,
                              after_alloc=(lambda m: comm_init(m, dist, rank, world)) if world > 1 else None,
                              n_cpu=world, my_rank=rank, device=local)
    f0, g0 = mg.stat("fast_launches"), mg.stat("general_launches")
    hist = [mg.mg_fas_fmg(False)] + [mg.mg_fas_vcycle() for _ in range(3)]
    fast, general = mg.stat("fast_launches") - f0, mg.stat("general_launches") - g0
    ck = phi_checksum(mg, P)
    fused, puts, fusedp = mg.stat("fused_smooths"), mg.stat("puts"), mg.stat("fused_prolongs")
    if world > 1:
        t = torch.tensor([ck - (1 << 64) if ck >= (1 << 63) else ck], dtype=torch.int64, device="cuda")
        dist.all_reduce(t)            # two's-complement wrap-around = sum mod 2^64
        ck = int(t.item()) % (1 << 64)
    mg.mg_deallocate_storage()
    out = {"n_ranks": world, "problem": "test_uniform_grid %d^3 box %d, FMG + 3 V(2,2) GSRB cycles" % (n, box),
           "fast_launches": fast, "general_launches": general, "fused_smooths": fused, "fused_prolongs": fusedp, "put_phases": puts}
    if rank == 0:
        from oracle import oracle as O
        O.lib().orc_set_omp_threads(host_threads())
        orc, _ = P.uniform_problem(O.OracleMg, n, box, load_balance="parents", n_cpu=host_threads())
        h_o = [orc.mg_fas_fmg(False)] + [orc.mg_fas_vcycle() for _ in range(3)]
        ck_o = phi_checksum(orc, P)
        out.update({"residuals": [float(x).hex() for x in hist], "oracle_residuals": [float(x).hex() for x in h_o],
                    "phi_checksum": "%016x" % ck, "oracle_phi_checksum": "%016x" % ck_o,
                    "bitwise": bool(h_o == hist and ck == ck_o)})
    return out


def build_performance_mg(om, P, args, dims, op, rank, world, local, dist, collapse):
    """tests/test_performance.f90:51-66 without host cc arrays (the run is device-resident)"""
    mg = om.Mg(op, P.mg_smoother_gs if getattr(args, "smoother", "gsrb") == "gs" else P.mg_smoother_gsrb, n_extra_vars=P.n_eps(op), n_cpu=world, my_rank=rank, device=local, gpu_resident=True)
    mg.set_int("host_cc", 0)
    if op in (P.mg_helmholtz, P.mg_vhelmholtz, P.mg_ahelmholtz):
        mg.set_lambda(1.0)
    if world > 1:
        mg.set_int("coarse_collapse_boxes", collapse)
    for nb in range(1, 7):
        if args.op == "laplacian" or (args.op == "vlaplacian" and nb == 1):
            mg.set_bc(nb, P.mg_iphi, P.mg_bc_dirichlet, 0.0)
        else:
            mg.set_bc(nb, P.mg_iphi, P.mg_bc_neumann, 0.0)
    mg.mg_set_methods()
    mg.mg_build_rectangle(dims, args.box, [1.0 / dims[0]] * 3, [0.0] * 3, [False] * 3)
    mg.mg_load_balance()
    mg.mg_load_balance_parents()
    mg.mg_allocate_storage()
    if world > 1:
        comm_init(mg, dist, rank, world)
    G, gpu = mg.G, mg.gpu
    # device-resident initial state: rhs = 1 on the interior of every level, phi = 0 (no host transfer needed)
    mg._gck(G.omg_set_var_const(gpu, P.mg_irhs, -20, 20, 1.0, 1))
    mg._gck(G.omg_set_var_const(gpu, P.mg_iphi, -20, 20, 0.0, 0))
    for d in range(P.n_eps(op)):
        mg._gck(G.omg_set_var_const(gpu, P.mg_iveps + d, -20, 20, 1.0, 0))     # eps = 1 incl. ghost faces
    if args.tail_cells >= 0:
        mg.set_option("tail_cells", args.tail_cells)
    if args.no_graph:
        mg.set_option("use_graph", 0)
    if args.no_fast:
        mg.set_option("fast_kernels", 0)
    # resident mode, and the host of this benchmark never reads mg_ires: the residual of a level is consumed in
    # registers by update_coarse (restricted at once) and is not written back (8 B per cell less); --store-res keeps
    # the reference's memory image of mg_ires
    mg.set_option("store_fine_residual", 1 if args.store_res else 0)
    for kv in args.opt:
        k, v = kv.split("=")
        mg.set_option(k, int(v))
    return mg


def timed_cycles(mg, args, barrier, dist, steps, warmup):
    """W warm-up cycles, then K cycles bracketed by barrier + synchronize; CUDA events on the handle's stream; max over ranks"""
    import torch
    mg.mg_fas_vcycle(want_res=args.max_res)               # uploads parameters; first cycle also builds the graph
    for _ in range(max(warmup - 1, 0)):
        mg.fas_vcycles(1, want_res=args.max_res)
    barrier()
    res, ms = mg.fas_vcycles(steps, want_res=args.max_res)
    barrier()
    if dist is not None:
        tms = torch.tensor([ms], device="cuda")
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        ms = float(tms.item())
    return res, ms


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--n", "--cells", dest="n", type=int, default=512, help="cells per direction (per GPU, or in total with --strong); "
                    "use --cells under torchrun, whose own parser trips over the abbreviation --n")
    ap.add_argument("--box", type=int, default=16)
    ap.add_argument("--max-res", action="store_true", help="request max_res in every timed cycle (extra residual pass over all levels)")
    ap.add_argument("--collapse", type=int, default=-1, help="coarse levels with fewer boxes per rank than this live on one GPU (N > 1; -1: default)")
    ap.add_argument("--tail-cells", type=int, default=-1, help="levels up to this many cells run inside the single-launch coarse tail")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-strong", action="store_true", help="skip the 1024^3 strong-scaling record")
    ap.add_argument("--no-parity", action="store_true", help="skip the bitwise parity check of the timed data plane")
    ap.add_argument("--no-configs", action="store_true", help="skip the record of BASELINE configs[3] (vlaplacian / Helmholtz 512^3, Neumann BC)")
    ap.add_argument("--no-amr", action="store_true", help="skip the record of BASELINE configs[2] (AMR grid of tests/test_refinement.f90, scaled up)")
    ap.add_argument("--strong-cells", type=int, default=1024, help="total domain of the strong-scaling record (BASELINE configs[4])")
    ap.add_argument("--parity-cells", type=int, default=256, help="total domain of the parity check")
    ap.add_argument("--cpu-budget", type=float, default=150.0, help="seconds the CPU arm may take (it cuts its cycle count, then its domain, to fit)")
    ap.add_argument("--e2e-cc", action="store_true", help="end-to-end leg moves cc slabs with ghost shells (omg_push/pull_var_packed) "
                    "instead of interior cells (omg_copy_to/from_tree)")
    ap.add_argument("--no-graph", action="store_true", help="launch kernels directly instead of replaying a CUDA graph")
    ap.add_argument("--no-fast", action="store_true", help="general-path kernels only")
    ap.add_argument("--store-res", action="store_true", help="write the residual of every level to mg_ires as the reference does "
                    "(default: resident mode without it -- nothing on the device reads it and this host never pulls it)")
    ap.add_argument("--smoother", default="gsrb", choices=["gsrb", "gs"], help="mg_smoother_gsrb (test_performance.f90, the headline) or the "
                    "reference's default mg_smoother_gs (lexicographic Gauss-Seidel inside each box); a side measurement, not the metric")
    ap.add_argument("--op", default="laplacian", choices=["laplacian", "vlaplacian", "helmholtz", "vhelmholtz", "ahelmholtz"],
                    help="operator (BASELINE.json configs[3]: vlaplacian / helmholtz with Neumann BC); default = the headline Laplacian")
    ap.add_argument("--strong", action="store_true", help="N > 1: keep the TOTAL domain at n^3 instead of n^3 per GPU for the headline value")
    ap.add_argument("--opt", action="append", default=[], help="library option key=value (omg_set_option), repeatable")
    ap.add_argument("--trace", default=None, metavar="PREFIX",
                    help="after the measurement: run 2 V-cycles with option trace and write the per-step timeline of every rank "
                         "to PREFIX.rank<r>.json (diagnostic; not part of any reported number)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dims = [args.n] * 3 if args.strong else domain_for(args.n, world)
    config = {"workload": workload_text(dims, args.box, args.max_res),
              "n_cycle_down": 2, "n_cycle_up": 2, "smoother": args.smoother, "box_size": args.box, "domain": dims,
              "partition": "mg_load_balance (Morton-contiguous), %d rank(s)" % world,
              "l2": "inputs larger than L2 (4 x %.2f GB per GPU vs 126 MB)" % (dims[0] * dims[1] * dims[2] / world * 8 / 1e9)}

    if args.impl == "reference":
        if rank != 0:
            return
        cb = cpu_reference_line(dims, args.box, args.steps, args.warmup, args.max_res, args.cpu_budget)
        ran = cb["ran"]
        # the line describes what RAN: when the budget cut the domain or the cycle count, config / steps / warmup say so
        if ran["domain"] != dims:
            config["sample_of"] = config["workload"]
            config["workload"] = workload_text(ran["domain"], args.box, args.max_res)
            config["domain"] = ran["domain"]
        config["partition"] = "%d emulated MPI ranks (OpenMP threads), mg_load_balance" % cb["cores"]
        config["l2"] = "host arm"
        line = {"impl": "reference", "metric": "3D Poisson FAS V-cycle throughput", "value": cb["value"], "unit": "GDOF/s",
                "n_gpus": args.gpus, "steps": ran["steps"], "warmup": ran["warmup"], "requested": {"steps": args.steps, "warmup": args.warmup},
                "ms_per_step": cb["ms_per_step"],
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": config, "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
                "e2e": {"value": cb["value"], "unit": "GDOF/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "ref_style_unknowns_per_us": cb["ref_style_unknowns_per_us"],
                "note": "reference MPI Fortran build unavailable on this box (no Fortran compiler / MPI, profiles/fortran_probe_r02.txt); CPU arm = oracle port"}
        print(json.dumps(line))
        return

    import torch
    import octree_mg_b200 as om
    import problems as P
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    op = {"laplacian": P.mg_laplacian, "vlaplacian": P.mg_vlaplacian, "helmholtz": P.mg_helmholtz, "vhelmholtz": P.mg_vhelmholtz,
          "ahelmholtz": P.mg_ahelmholtz}[args.op]
    if args.op != "laplacian":
        config["workload"] = config["workload"].replace("7-point Laplacian", "%s (eps = 1, lambda = 1), Neumann-0 BC%s" % (
            args.op, " + Dirichlet on -x" if args.op == "vlaplacian" else ""))
    collapse = args.collapse if args.collapse >= 0 else 128
    if world > 1:
        config["coarse_collapse_boxes_per_rank"] = collapse
    config["options"] = {"store_fine_residual": 1 if args.store_res else 0}
    for kv in args.opt:
        k, v = kv.split("=")
        config["options"][k] = int(v)
    mg = build_performance_mg(om, P, args, dims, op, rank, world, local, dist, collapse)
    dof = dims[0] * dims[1] * dims[2]
    G, gpu = mg.G, mg.gpu

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    phase("handle built")
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    res, ms = timed_cycles(mg, args, barrier, dist, args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    l0, f0, g0 = mg.stat("launches"), mg.stat("fast_launches"), mg.stat("general_launches")
    mg.fas_vcycles(1, want_res=args.max_res)
    launches_per_step = mg.stat("launches") - l0
    fast_per_step, general_per_step = mg.stat("fast_launches") - f0, mg.stat("general_launches") - g0
    res_after, _ = mg.fas_vcycles(1, want_res=True)      # untimed: residual after warm-up + timed + 2 cycles
    n_boxes_total = mg.n_boxes

    phase("timed cycles done: %.3f ms per cycle" % (ms / args.steps))
    # --- roofline of the dominant kernel (finest-level GSRB colour half-sweep), timed live with CUDA events
    mg.set_option("profile", 1)
    mg.fas_vcycles(2)
    sm_ms = float(G.omg_last_ms(gpu, b"smooth"))
    n_sm = int(G.omg_last_ms(gpu, b"n_smooth"))
    prof = {k: float(G.omg_last_ms(gpu, k.encode())) for k in ("smooth", "resid_restrict", "corr_prolong", "resid")}
    mg.set_option("profile", 0)
    timers = mg.mg_timers()
    if args.trace:
        mg.set_option("trace", 1)
        mg.fas_vcycles(2)
        steps_tr = mg.trace()
        mg.set_option("trace", 0)
        by = {}
        for name, lvl, a, b, t in steps_tr:
            k = "%s@%d" % (name, lvl)
            by[k] = [by.get(k, [0.0, 0])[0] + t, by.get(k, [0.0, 0])[1] + 1]
        with open("%s.rank%d.json" % (args.trace, rank), "w") as f:
            json.dump({"cycles": 2, "total_ms": sum(t for *_, t in steps_tr),
                       "by_step_ms": {k: [round(v[0], 4), v[1]] for k, v in sorted(by.items(), key=lambda kv: -kv[1][0])},
                       "steps": [[n_, l_, a_, b_, round(t_, 4)] for n_, l_, a_, b_, t_ in steps_tr]}, f)
    peak, peak_src = peaks()
    traffic, traffic_src = ncu_traffic(args.n, args.box)
    cells_per_rank = dof / world
    achieved = SMOOTH_BYTES_PER_CELL * cells_per_rank / (sm_ms * 1e-3) / 1e9 if sm_ms > 0 else None
    roofline = {"bound": "hbm", "kernel": "gsrb colour half-sweep, finest level (%d launches timed)" % n_sm,
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": (achieved / peak) if achieved else None,
                "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "algorithmic_bytes_per_launch": SMOOTH_BYTES_PER_CELL * cells_per_rank,
                "ms_per_launch": sm_ms, "finest_level_ms": prof}
    value = dof * args.steps / (ms * 1e-3) / 1e9
    cycle_frac = (BYTES_PER_DOF_VCYCLE + (BYTES_PER_DOF_MAXRES if args.max_res else 0.0)) * value / peak / world
    # mg_timers_show (src/m_data_structures.f90:562-581): min / max over ranks of the accumulated timers
    timers_minmax = None
    if dist is not None:
        names = sorted(timers)
        tv = torch.tensor([timers[k] for k in names], dtype=torch.float64, device="cuda")
        tmin, tmax = tv.clone(), tv.clone()
        dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        timers_minmax = {k: [float(a), float(b)] for k, a, b in zip(names, tmin.tolist(), tmax.tolist())}
    else:
        timers_minmax = {k: [v, v] for k, v in timers.items()}

    # --- end to end through the C ABI with host buffers: the application's phi and rhs (pinned host arrays, interior
    # cells of the grid leaves, as mg_copy_to_tree of coupling_amrvac/mod_multigrid_coupling.t:133-171 takes them) go to
    # the device tree, one V-cycle runs, phi comes back (mg_copy_from_tree :206-234).  Leaves carry every input:
    # mg_fas_vcycle overwrites phi and rhs of all parent boxes from the finer level before reading them (update_coarse,
    # src/m_multigrid.f90:347-384), and phi's ghost cells are derived on the device.  --e2e-cc moves the reference's
    # cc slabs (ghost shells included) of the leaf level instead.
    phase("profile / trace done")
    e2e = None
    if not args.no_e2e:
        top = mg.highest_lvl
        if args.e2e_cc:
            nd = int(G.omg_packed_size(gpu, top, top))
        else:
            nd = int(G.omg_copy_size(gpu, -20, 20, 1, 0))
        hphi = torch.zeros(nd, dtype=torch.float64).pin_memory()
        hrhs = torch.ones(nd, dtype=torch.float64).pin_memory()
        r = C.c_double(0.0)
        k2 = max(1, min(args.steps, 3))

        def step():
            if args.e2e_cc:
                mg._gck(G.omg_push_var_packed(gpu, P.mg_iphi, top, top, C.c_void_p(hphi.data_ptr()), nd))
                mg._gck(G.omg_push_var_packed(gpu, P.mg_irhs, top, top, C.c_void_p(hrhs.data_ptr()), nd))
            else:
                mg._gck(G.omg_copy_to_tree(gpu, P.mg_iphi, -20, 20, 1, 0, C.c_void_p(hphi.data_ptr()), nd))
                mg._gck(G.omg_copy_to_tree(gpu, P.mg_irhs, -20, 20, 1, 0, C.c_void_p(hrhs.data_ptr()), nd))
            mg._gck(G.omg_fas_vcycle(gpu, 0, 0, 1, int(args.max_res), C.byref(r)))
            if args.e2e_cc:
                mg._gck(G.omg_pull_var_packed(gpu, P.mg_iphi, top, top, C.c_void_p(hphi.data_ptr()), nd, 1))
            else:
                mg._gck(G.omg_copy_from_tree(gpu, P.mg_iphi, -20, 20, 1, 0, C.c_void_p(hphi.data_ptr()), nd))
        step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(k2):
            step()
        barrier()
        dt = time.perf_counter() - t0
        if dist is not None:
            tt = torch.tensor([dt], device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dt = float(tt.item())
        e2e = {"value": dof * k2 / dt / 1e9, "unit": "GDOF/s", "h2d_bytes_per_step": 2 * nd * 8 * world,
               "d2h_bytes_per_step": nd * 8 * world + 8, "steps": k2, "ms_per_step": dt / k2 * 1e3,
               "pcie_gbs_per_gpu": 3 * nd * 8 / (dt / k2) / 1e9,
               "what": ("omg_push_var_packed(phi, rhs) of the leaf level from pinned host slabs in the reference's cc layout (ghost shells included) + omg_fas_vcycle + omg_pull_var_packed(phi with ghost cells)"
                        if args.e2e_cc else "omg_copy_to_tree(phi, rhs: interior cells of the leaves, pinned host arrays) + omg_fas_vcycle + omg_copy_from_tree(phi)")}
        del hphi, hrhs

    # mg_deallocate_storage on every rank (ordered release of the peer mappings), then the follow-up problems are
    # created on the same process group: create / destroy / create of the multi-rank handle is part of every run
    phase("end-to-end leg done; deallocating")
    mg.mg_deallocate_storage()
    del mg
    phase("deallocated")
    barrier()

    # --- strong scaling of BASELINE configs[4]: the same 1024^3 domain on the N GPUs of this run
    strong = None
    if not args.no_strong and args.op == "laplacian":
        sdims = [args.strong_cells] * 3
        if sdims == dims:
            strong = {"domain": sdims, "n_gpus": world, "ms_per_step": ms / args.steps, "value": value, "unit": "GDOF/s",
                      "steps": args.steps, "warmup": args.warmup, "same_run_as_value": True}
        else:
            smg = build_performance_mg(om, P, args, sdims, op, rank, world, local, dist, collapse)
            ks = max(3, min(args.steps, 10))
            _, sms = timed_cycles(smg, args, barrier, dist, ks, 3)
            sres, _ = smg.fas_vcycles(1, want_res=True)
            strong = {"domain": sdims, "n_gpus": world, "ms_per_step": sms / ks, "value": sdims[0] ** 3 * ks / (sms * 1e-3) / 1e9,
                      "unit": "GDOF/s", "steps": ks, "warmup": 3, "max_res_after": sres[0], "launches_per_step": None}
            l0 = smg.stat("launches")
            smg.fas_vcycles(1, want_res=args.max_res)
            strong["launches_per_step"] = smg.stat("launches") - l0
            phase("strong record done: %.3f ms per cycle; deallocating" % (sms / ks))
            smg.mg_deallocate_storage()
            del smg
            barrier()

    # --- bitwise parity of the data plane that was just timed
    parity = None
    if not args.no_parity and args.op == "laplacian":
        phase("parity check")
        parity = parity_check(om, P, args.parity_cells, args.box, rank, world, local, dist, collapse)
        phase("parity check done")

    # --- BASELINE configs[3]: variable-coefficient Laplacian and Helmholtz, 512^3, Neumann BC: V(2,2) cycles and one FMG cycle
    other = None
    if world == 1 and not args.no_configs and args.op == "laplacian" and not args.no_strong:
        other = {}
        for name, oc in (("vlaplacian", P.mg_vlaplacian), ("helmholtz", P.mg_helmholtz)):
            phase("configs[3]: %s" % name)
            a2 = argparse.Namespace(**vars(args)); a2.op = name
            omg2 = build_performance_mg(om, P, a2, dims, oc, rank, world, local, dist, collapse)
            k3 = max(3, min(args.steps, 10))
            _, ms3 = timed_cycles(omg2, a2, barrier, dist, k3, 3)
            r3, _ = omg2.fas_vcycles(1, want_res=True)
            omg2._gck(omg2.G.omg_set_var_const(omg2.gpu, P.mg_iphi, -20, 20, 0.0, 0))
            omg2.mg_fas_fmg(False)                                   # builds the FMG program and its graph
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            rf = omg2.mg_fas_fmg(False)                              # FMG solve from phi = 0 (returns max_res: synchronises)
            t_fmg = (time.perf_counter() - t0) * 1e3
            other[name + "_%d" % dims[0]] = {"ms_per_vcycle": ms3 / k3, "value": dof * k3 / (ms3 * 1e-3) / 1e9, "unit": "GDOF/s", "steps": k3,
                                           "max_res_after": r3[0], "fmg_ms": t_fmg, "fmg_max_res": rf,
                                           "bc": "Neumann 0" + (" + Dirichlet on -x" if name == "vlaplacian" else ""), "eps": 1.0, "lambda": 1.0 if name == "helmholtz" else None}
            omg2.mg_deallocate_storage()
            del omg2
    # --- BASELINE configs[2]: tests/test_refinement.f90 scaled up: 256^3 base grid + 3 levels of centred refinement
    # (4096 boxes of 16^3 on each of the 4 levels, refinement-boundary ghost interpolation on every fine level), Dirichlet 0
    if other is not None and not args.no_amr:
        phase("configs[2]: AMR")
        mg3 = om.Mg(P.mg_laplacian, P.mg_smoother_gsrb, device=local, gpu_resident=True)
        mg3.set_int("host_cc", 0)
        for nb in range(1, 7):
            mg3.set_bc(nb, P.mg_iphi, P.mg_bc_dirichlet, 0.0)
        mg3.mg_set_methods()
        P.build_amr_tree(mg3, 4, [256] * 3, args.box, [1.0 / 256] * 3)
        mg3._gck(mg3.G.omg_set_var_const(mg3.gpu, P.mg_irhs, -20, 20, 1.0, 1))
        mg3._gck(mg3.G.omg_set_var_const(mg3.gpu, P.mg_iphi, -20, 20, 0.0, 0))
        mg3.set_option("store_fine_residual", 1 if args.store_res else 0)
        unknowns = int(mg3.mg_number_of_unknowns())
        cells_all = sum(len(mg3.ids(l)) for l in range(1, mg3.highest_lvl + 1)) * args.box ** 3
        k2 = max(3, min(args.steps, 10))
        l0, f0 = mg3.stat("launches"), mg3.stat("fast_launches")
        _, ms2 = timed_cycles(mg3, args, barrier, dist, k2, 3)
        r2, _ = mg3.fas_vcycles(1, want_res=True)
        other["refinement_256_4lvl"] = {"ms_per_vcycle": ms2 / k2, "value": unknowns * k2 / (ms2 * 1e-3) / 1e9, "unit": "GDOF/s (leaf cells)",
                                        "unknowns": unknowns, "cells_on_levels_1_to_4": cells_all,
                                        "gcells_per_s_all_levels": cells_all * k2 / (ms2 * 1e-3) / 1e9, "steps": k2, "max_res_after": r2[0],
                                        "boxes_per_level": [len(mg3.ids(l)) for l in range(mg3.lowest_lvl, mg3.highest_lvl + 1)],
                                        "launches": mg3.stat("launches") - l0, "fast_launches": mg3.stat("fast_launches") - f0}
        mg3.mg_deallocate_storage()
        del mg3
    if rank != 0:
        finish(world, dist)
        return
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        cb = cpu_reference_line(dims, args.box, 3, 1, args.max_res, 60.0)
        cpu = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
    line = {"metric": "3D Poisson FAS V-cycle throughput", "value": value, "unit": "GDOF/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "strong" if args.strong else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
            "roofline": roofline, "cycle_hbm_frac": cycle_frac, "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": launches_per_step * args.steps if launches_per_step else None,
            "launches_per_step": launches_per_step, "fast_launches_per_step": fast_per_step, "general_launches_per_step": general_per_step,
            "clocks": clocks, "residuals": res, "max_res_after": res_after,
            # tests/test_performance.f90:90-92: n_boxes * box_size^3 (all boxes on all levels) per microsecond
            "ref_style_unknowns_per_us": n_boxes_total * args.box ** 3 / (ms / args.steps * 1e3),
            "timers_min_max_s": timers_minmax, "strong": strong, "parity_check": parity, "configs": other}
    print(json.dumps(line))
    finish(world, dist)
    if parity is not None and not parity.get("bitwise", False):
        sys.stderr.write("bench.py: parity check FAILED: %r\n" % parity)
        sys.exit(4)


_T0 = time.time()


def phase(msg):
    """progress line on stderr (every rank): where a multi-rank run stands; the last line tells where a blocked run stopped"""
    if int(os.environ.get("WORLD_SIZE", "1")) > 1 or os.environ.get("OMG_BENCH_VERBOSE"):
        sys.stderr.write("[bench rank %s %.1fs] %s\n" % (os.environ.get("RANK", "0"), time.time() - _T0, msg))
        sys.stderr.flush()


def finish(world, dist):
    """Multi-rank runs: the library handles are already destroyed (ordered release, omg_comm.cu); destroy the process
    group and leave through the interpreter.  A watchdog turns a teardown that blocks (round 1: runs that printed their
    line and never exited) into a plain exit with a message instead of a hung box."""
    sys.stdout.flush()
    sys.stderr.flush()
    if world == 1:
        return
    if os.environ.get("OMG_BENCH_PLAIN_EXIT") == "0":
        os._exit(0)
    wd = threading.Timer(60.0, lambda: (sys.stderr.write("bench.py: teardown did not finish within 60 s; leaving through os._exit\n"),
                                        sys.stderr.flush(), os._exit(0)))
    wd.daemon = True
    wd.start()
    try:
        dist.barrier()
        dist.destroy_process_group()
        sys.stderr.write("bench.py: rank %s teardown clean\n" % os.environ.get("RANK", "0"))
    except Exception as e:      # noqa: BLE001
        sys.stderr.write("bench.py: teardown raised %r\n" % (e,))
    sys.stderr.flush()


if __name__ == "__main__":
    # a run that blocks (a peer that never arrives, a teardown that never returns) must not hold the box: leave after
    # 25 minutes with a non-zero status instead
    _watchdog = threading.Timer(1500.0, lambda: (sys.stderr.write("bench.py: watchdog expired\n"), sys.stderr.flush(), os._exit(3)))
    _watchdog.daemon = True
    _watchdog.start()
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import faulthandler
        faulthandler.dump_traceback_later(180, repeat=True, file=sys.stderr)     # a blocked rank shows where it stands
    main()
