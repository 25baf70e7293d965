# C5 with the x gathers evict-last (kernel 5) and an L2 set-aside for persisting lines: does x stay in L2 against the streams?
cd /root/repo 2>/dev/null || cd "$GRAFT_REPO_ROOT"
out=gpurun_out/r2_exp_l2_persist.txt; : > $out
for k in 0 5; do for mb in -1 0 16 64; do
  echo "== csr-kernel $k l2-persist-mb $mb" >> $out
  python bench.py --workload c5 --csr-kernel $k --l2-persist-mb $mb --extras 0 --no-cpu --no-e2e --steps 20 --warmup 3 2> gpurun_out/exp_l2.err | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms %.4f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))" >> $out
  grep "l2 persisting" gpurun_out/exp_l2.err >> $out
done; done
