# C5 / C1 with the bulk-copy-fed stream kernel (tuning build): CME213_CSR_CFG=threads,tile,7
cd /root/repo 2>/dev/null || cd "$GRAFT_REPO_ROOT"
out=gpurun_out/r2_exp_csr_bulk.txt; : > $out
CME213_CSR_CFG=128,1024,7 timeout 300 python -m pytest tests/test_spmv_gpu.py -m gpu -q -x -k "not pagerank and not graph" 2>&1 | tail -4 >> $out
for cfg in ${CFGS:-128,1024,0 128,1024,7 64,512,7 256,2048,7 128,1536,7 128,2048,7 64,1024,7}; do
  echo "== c5 CME213_CSR_CFG=$cfg" >> $out
  CME213_CSR_CFG=$cfg python bench.py --workload c5 --csr-kernel 3 --steps 20 --warmup 3 --extras 0 --no-cpu --no-e2e 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms %.4f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))" >> $out
done
for cfg in 256,2048,0 256,2048,7 128,1024,7; do
  echo "== c1 CME213_CSR_CFG=$cfg" >> $out
  CME213_CSR_CFG=$cfg python bench.py --workload c1 --csr-kernel 3 --steps 50 --warmup 5 --extras 0 --no-cpu --no-e2e 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms/pass %.5f frac %.3f' % (d['ms_per_step']/8, d['roofline']['frac']))" >> $out
done
