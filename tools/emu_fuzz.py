"""Randomised sweep of the CTA emulator comparisons (tests/device_host/fast_emu.cpp): random operator, box size, brick shape,
boundary types and values, boundary representation, rank count and seed, for a time budget in seconds (default 420).
Run from the repo root after `python __graft_entry__.py`.  Last run: 1406 cases, 0 mismatches, 0 violations."""
import sys
import ctypes as C, itertools, random, time
L = C.CDLL("tests/device_host/libfastemu.so")
six = lambda t, v: (t * 6)(*v)
for f in (L.fe_one_level, L.fe_two_levels):
    f.argtypes = [C.c_int] * 5 + [C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_double, C.c_ulonglong]
L.fe_ranks.argtypes = [C.c_int] * 7 + [C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_double, C.c_ulonglong]
L.fe_tail.argtypes = [C.c_int] * 8 + [C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_double, C.c_ulonglong]
L.dh_set_planes.argtypes = [C.c_int]
BUDGET = float(sys.argv[1]) if len(sys.argv) > 1 else 420.0
rnd = random.Random(int(sys.argv[2]) if len(sys.argv) > 2 else 2024)
t0 = time.time(); n = 0; bad = []
while time.time() - t0 < BUDGET:
    types = [rnd.choice([-10, -11, -12]) for _ in range(6)]
    vals = [rnd.uniform(-2, 2) for _ in range(6)]
    op, lam = rnd.choice([(1, 0.0), (2, 0.0), (3, rnd.uniform(0.1, 5)), (4, rnd.uniform(0.1, 5))])
    nc = rnd.choice([16, 16, 32])
    shape = rnd.choice([(1, 1, 1), (1, 1, 3), (1, 3, 1), (3, 1, 1), (2, 2, 2), (1, 2, 3), (2, 3, 1)])
    planes = rnd.choice([0, 1, 2])
    seed = rnd.randrange(1, 10**6)
    L.dh_set_planes(planes)
    kind = rnd.choice(["one", "two", "ranks", "tail"])
    if kind == "one":
        r = L.fe_one_level(op, nc, *shape, six(C.c_int, types), six(C.c_double, vals), lam, seed)
    elif kind == "two":
        sh = rnd.choice([(1, 1, 1), (2, 1, 1), (1, 1, 2)])
        r = L.fe_two_levels(op, nc, *sh, six(C.c_int, types), six(C.c_double, vals), lam, seed)
    elif kind == "ranks":
        L.dh_set_planes(0)
        nb = shape[0] * shape[1] * shape[2]
        if nb < 2: continue
        nr = rnd.choice([k for k in (2, 3, 4) if k <= nb])
        r = L.fe_ranks(op, nc, *shape, nr, rnd.choice([0, 1]), six(C.c_int, types), six(C.c_double, vals), lam, seed)
    else:
        if planes == 2: L.dh_set_planes(1)
        opt = rnd.choice([1, 2, 3, 4, 5])
        r = L.fe_tail(opt, 256, 3, *rnd.choice([(1, 1, 1), (2, 1, 1), (1, 2, 2)]), rnd.choice([0, 1]), rnd.choice([0, 1]), six(C.c_int, types), six(C.c_double, vals), lam, seed)
    n += 1
    if r != 0: bad.append((kind, op, nc, shape, planes, types, seed, r)); print("MISMATCH", bad[-1], flush=True)
L.dh_set_planes(0)
print("cases", n, "mismatches", len(bad), "violations", L.fe_violations())
