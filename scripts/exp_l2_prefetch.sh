# NOTE: the K = 3..6 variants this script drives were removed after the measurement (no effect; profiles/r02_exp_l2_prefetch_hints.txt,
# git log -S'LTC128B' finds the commit); kept as the record of how the numbers were taken.
# C5 (tuning build): x gathers / matrix streams with L2 prefetch-size hints, and the device's L2 fetch granularity
# CME213_CSR_CFG=128,1024,K: K = 0 plain, 3 gathers .L2::128B, 4 gathers .L2::256B, 5 streams .L2::256B, 6 = 3 + 5
cd /root/repo 2>/dev/null || cd "$GRAFT_REPO_ROOT"
out=gpurun_out/r2_exp_l2_prefetch.txt; : > $out
one() { echo "== CFG=$1 fetch-gran=$2" >> $out
  CME213_CSR_CFG=$1 python bench.py --workload c5 --csr-kernel 3 --l2-fetch-gran $2 --extras 0 --no-cpu --no-e2e --steps 20 --warmup 3 2> gpurun_out/exp_l2.err | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms %.4f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))" >> $out
  grep "l2 max fetch" gpurun_out/exp_l2.err >> $out; }
for k in 0 3 4 5 6; do one 128,1024,$k 0; done
one 128,1024,0 32; one 128,1024,0 128; one 128,1024,6 128
