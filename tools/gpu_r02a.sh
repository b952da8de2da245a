#!/bin/bash
# round-2 pass a (1 GPU): everything written after round 1's GPU budget ran out + the round-2 parity work.
#   probe of the box (Fortran / MPI, host cores, memory, topology) -> profiles/fortran_probe_r02.txt
#   the whole GPU suite (no opt-in gates left), smoke(), default bench line (strong record, parity check, CPU arm),
#   reference arm as the driver calls it, option sweeps, launch list, full ncu captures of the three finest-level kernels
set -x
mkdir -p gpurun_out
{
  echo "# probe of the GPU box, $(date -u +%FT%TZ)"
  for c in gfortran mpif90 mpifort nvfortran flang ifort ifx mpirun mpiexec f951; do printf '%-10s ' $c; command -v $c || echo "(absent)"; done
  find / -name 'f951' -o -name 'libmpi.so*' 2>/dev/null | head -5
  echo "# host"; nproc; grep -c ^processor /proc/cpuinfo; grep MemAvailable /proc/meminfo; lscpu | grep -i 'model name\|numa\|socket' | head -8
  echo "# affinity"; python -c "import os; print(sorted(os.sched_getaffinity(0))[:4], len(os.sched_getaffinity(0)))"
  nvidia-smi -L; nvidia-smi topo -m 2>/dev/null | head -14
} > gpurun_out/fortran_probe_r02.txt 2>&1
( time timeout 1700 python -m pytest tests -m gpu -q --durations=8 ) > gpurun_out/pytest_r02a.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_r02a.log
tail -22 gpurun_out/pytest_r02a.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_r02a.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke_r02a.log
( time timeout 900 python bench.py ) > gpurun_out/bench_r02a.json 2> gpurun_out/bench_r02a.err; echo "bench rc=$?"; tail -5 gpurun_out/bench_r02a.err
( time timeout 600 python bench.py --impl reference --steps 20 --warmup 5 ) > gpurun_out/bench_r02a_reference.json 2> gpurun_out/bench_r02a_reference.err; echo "ref rc=$?"
Q="--no-cpu-baseline --no-e2e --no-strong --no-parity"
timeout 300 python bench.py $Q --opt pdl=1 > gpurun_out/bench_r02a_pdl.json 2> gpurun_out/bench_r02a_pdl.err
timeout 300 python bench.py $Q --opt tail_cells=32768 > gpurun_out/bench_r02a_tc32k.json 2> gpurun_out/bench_r02a_tc32k.err
timeout 300 python bench.py $Q --opt tail_cells=32768 --opt pdl=1 > gpurun_out/bench_r02a_tc32k_pdl.json 2> gpurun_out/bench_r02a_tc32k_pdl.err
for OP in helmholtz vlaplacian vhelmholtz ahelmholtz; do
  timeout 300 python bench.py $Q --op $OP > gpurun_out/bench_r02a_$OP.json 2> gpurun_out/bench_r02a_$OP.err
done
timeout 300 python bench.py $Q --op vlaplacian --opt tail_simple=2 > gpurun_out/bench_r02a_vlaplacian_ts2.json 2> gpurun_out/bench_r02a_vlaplacian_ts2.err
CMD="python bench.py --steps 2 --warmup 3 $Q"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r02a.csv $CMD > gpurun_out/ncu_list_r02a.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_smooth_lpl_fast -c 1 -o gpurun_out/prof_r02a_smooth $CMD > gpurun_out/ncu_r02a_1.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_stencil_fast -c 2 -o gpurun_out/prof_r02a_stencil $CMD > gpurun_out/ncu_r02a_2.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_corr_prolong_fast -s 3 -c 1 -o gpurun_out/prof_r02a_prolong $CMD > gpurun_out/ncu_r02a_3.log 2>&1
python - <<'PY'
import glob, json
for f in sorted(glob.glob("gpurun_out/bench_r02a*.json")):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, round(d["value"], 2), "GDOF/s", round(d["ms_per_step"], 3), "ms", "e2e", d.get("e2e") and round(d["e2e"]["value"], 2),
              "roof", d.get("roofline") and round(d["roofline"]["frac"] or 0, 3), "strong", d.get("strong") and round(d["strong"]["ms_per_step"], 2),
              "parity", d.get("parity_check") and d["parity_check"].get("bitwise"), "fast/gen", d.get("fast_launches_per_step"), d.get("general_launches_per_step"))
    except Exception as e:
        print(f, "FAILED", e)
PY
ls -la gpurun_out | tail -30
