# bash tools/ncu_quick.sh <out.csv> <prof_render args...>: a few issue / i-cache counters of one render_sorted launch (single ncu pass)
out=$1; shift
M=sm__icc_request_hit_rate.pct,gcc__cache_requests_type_instruction.sum,smsp__inst_executed.sum,gpu__time_duration.sum,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio,smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio,smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__thread_inst_executed_per_inst_executed.ratio,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum
timeout 600 ncu --metrics $M --clock-control none -k regex:render_sorted -s 1 -c 1 --csv --log-file $out python tools/prof_render.py "$@" > /dev/null 2>&1
python - "$out" <<'PY'
import csv,sys
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>10]
h=rows[0]; i=h.index("Metric Name"); v=h.index("Metric Value")
for r in rows[1:]: print(f"  {r[i]:90} {r[v]}")
PY
