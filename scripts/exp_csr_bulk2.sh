# NOTE: K = 8 (1536 resident threads) was removed after the measurement; K = 7 is kernel id CME213_CSR_STREAM_BULK now.  Record of how
# profiles/r02_exp_csr_bulk.txt was taken.
# second round: bulk kernel with 1536 resident threads (K = 8: all eight gathers of a thread in flight), and 2-D grids of several sizes
cd /root/repo 2>/dev/null || cd "$GRAFT_REPO_ROOT"
out=gpurun_out/r2_exp_csr_bulk2.txt; : > $out
CME213_CSR_CFG=128,1024,8 timeout 300 python -m pytest tests/test_spmv_gpu.py -m gpu -q -x -k "not pagerank and not graph" 2>&1 | tail -2 >> $out
for cfg in 128,1024,8 256,2048,8 64,512,8; do
  echo "== c5 CME213_CSR_CFG=$cfg" >> $out
  CME213_CSR_CFG=$cfg python bench.py --workload c5 --csr-kernel 3 --steps 20 --warmup 3 --extras 0 --no-cpu --no-e2e 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms %.4f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))" >> $out
done
for n in 512 1024 2048 4096; do for cfg in 256,2048,0 128,1024,7 128,1024,8 256,2048,7; do
  echo "== c1 n=$n CME213_CSR_CFG=$cfg" >> $out
  CME213_CSR_CFG=$cfg python bench.py --workload c1 --c1-n $n --csr-kernel 3 --steps 30 --warmup 5 --extras 0 --no-cpu --no-e2e 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms/pass %.5f frac %.3f' % (d['ms_per_step']/d['config']['passes_per_step'], d['roofline']['frac']))" >> $out
done; done
