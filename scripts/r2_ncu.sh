# ncu captures of round 2 (one GPU).  Every command runs once without ncu first (must exit 0), then under ncu.
cd /root/repo 2>/dev/null || cd "$GRAFT_REPO_ROOT"
cap() {  # name, kernel regex, skip, bench args...
  name=$1; regex=$2; skip=$3; shift 3
  python bench.py "$@" --extras 0 --no-cpu --no-e2e --steps 3 --warmup 3 > gpurun_out/r2_plain_$name.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$regex -s $skip -c 1 -f -o gpurun_out/r2_prof_$name \
      python bench.py "$@" --extras 0 --no-cpu --no-e2e --steps 3 --warmup 3 > gpurun_out/r2_ncu_$name.log 2>&1
  echo "$name rc=$?" >> gpurun_out/r2_ncu_status.txt
}
rm -f gpurun_out/r2_ncu_status.txt
cap c5 csr_stream_lean 4 --workload c5
cap pagerank csr_stream_lean 4 --workload pagerank
cap hyb_tail csr_stream_balanced 4 --workload hyb
cap hyb_ell ell_v4 4 --workload hyb
cap hyb_window_tail csr_stream_balanced 4 --workload hyb_window
cap hyb_coo_tail lookback_scan_kernel 4 --workload hyb_coo
cap csr_powerlaw csr_stream 4 --workload csr_powerlaw
cap csr_window csr_stream 4 --workload csr_window
cap fp_fused lookback_scan_kernel 4 --workload fp
cap fp_stream scan_bulk_blocked 8 --workload fp_iterate
cap c1 csr_stream_lean 10 --workload c1
cap segscan scan_bulk_blocked 4 --workload segscan
# launch list of the default command
python bench.py --steps 2 --warmup 3 --extras 0 --no-cpu --no-e2e > gpurun_out/r2_plain_list.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2_launches_c5.csv python bench.py --steps 2 --warmup 3 --extras 0 --no-cpu --no-e2e > gpurun_out/r2_ncu_list.log 2>&1
