#!/bin/bash
# round-2 pass zf (1 GPU): which pipe bounds the variable-coefficient kernels -- fp64 instruction counts and pipe utilisation of the
# vlaplacian smoother and residual kernels (bitwise and tolerance mode) next to DRAM throughput
set -x
mkdir -p gpurun_out
T=r02zf
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__inst_executed_pipe_fp64.sum,sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active,smsp__inst_executed_pipe_lsu.sum,smsp__inst_executed_pipe_alu.sum,smsp__inst_executed_pipe_fma.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,dram__bytes_read.sum,dram__bytes_write.sum,smsp__warp_issue_stalled_math_pipe_throttle_per_warp_active.pct,smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct,smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct,smsp__warp_issue_stalled_wait_per_warp_active.pct,smsp__warp_issue_stalled_barrier_per_warp_active.pct
Q="--op vlaplacian --no-cpu-baseline --no-amr --no-e2e --no-parity --no-strong --steps 2 --warmup 3"
for MODE in 0 1; do
  timeout 400 ncu --metrics $M --clock-control none -k regex:'k_smooth_lpl_fast|k_stencil_fast' --launch-skip 60 -c 24 --csv --log-file gpurun_out/ncu_vc_${T}_cm${MODE}.csv python bench.py $Q --opt coeff_math=$MODE > gpurun_out/ncu_vc_${T}_cm${MODE}.log 2>&1; echo "ncu rc=$?"
done
python - <<'PY'
import csv, collections
for mode in (0, 1):
    rows = list(csv.reader(open("gpurun_out/ncu_vc_r02zf_cm%d.csv" % mode)))
    hdr = next(r for r in rows if "Kernel Name" in r)
    ki, gi, mi, vi = hdr.index("Kernel Name"), hdr.index("Grid Size"), hdr.index("Metric Name"), hdr.index("Metric Value")
    acc = collections.OrderedDict()
    for r in rows:
        if len(r) != len(hdr) or not r[0].isdigit() or "32768" not in r[gi]:
            continue
        a = acc.setdefault((r[ki].split("(")[0][-48:], r[mi]), [])
        a.append(float(r[vi].replace(",", "")))
    print("# coeff_math =", mode, "(finest level, 32768 boxes; mean over the captured launches)")
    for (k, m), v in acc.items():
        print("%-50s %-75s %14.4g  (n=%d)" % (k, m, sum(v) / len(v), len(v)))
PY
