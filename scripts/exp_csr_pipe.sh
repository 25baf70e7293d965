# NOTE: the K = 9 / 10 hooks were removed when the two-stage kernel became kernel id CME213_CSR_STREAM_PIPE (7 | 1 << 8); the three-stage form
# lost and was not kept.  Record of how profiles/r02_exp_csr_pipe.txt was taken; scripts/exp_c1_kernels.py times the shipped ids.
# pipelined persistent bulk-copy-fed stream kernel (tuning build): CME213_CSR_CFG=128,1024,9 (2 stages) / ,10 (3 stages)
cd /root/repo 2>/dev/null || cd "$GRAFT_REPO_ROOT"
out=gpurun_out/r2_exp_csr_pipe.txt; : > $out
for k in 9 10; do
  CME213_CSR_CFG=128,1024,$k timeout 200 python -m pytest tests/test_spmv_gpu.py tests/test_full_size_gpu.py -m gpu -q -x -k "laplacian or c5_csr_dia or c1_csr" 2>&1 | tail -2 >> $out
done
for cfg in 128,1024,0 128,1024,9 128,1024,10; do
  echo "== c5 CME213_CSR_CFG=$cfg" >> $out
  CME213_CSR_CFG=$cfg timeout 120 python bench.py --workload c5 --csr-kernel 3 --steps 20 --warmup 3 --extras 0 --no-cpu --no-e2e 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms %.4f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))" >> $out
done
for cfg in 128,1024,7 128,1024,9 128,1024,10; do for n in 1024 4096; do
  echo "== c1 n=$n CME213_CSR_CFG=$cfg" >> $out
  CME213_CSR_CFG=$cfg timeout 120 python bench.py --workload c1 --c1-n $n --csr-kernel 3 --steps 30 --warmup 5 --extras 0 --no-cpu --no-e2e 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms/pass %.5f frac %.3f' % (d['ms_per_step']/d['config']['passes_per_step'], d['roofline']['frac']))" >> $out
done; done
