# stencil kernel choice (tuning build: python 2012-04_stanford_cme213_b200/build.py --tuning): CME213_STENCIL_G = 0 (one-vector kernel), 1, 2, 4 lines per thread
cd /root/repo 2>/dev/null || cd "$GRAFT_REPO_ROOT"
python -m pytest tests/test_dia_gpu.py -m gpu -x -q 2>&1 | tail -3 > gpurun_out/r2_stencil_sweep.txt
for g in 0 1 2 4; do
  for w in stencil c1; do :; done
  echo "== CME213_STENCIL_G=$g" >> gpurun_out/r2_stencil_sweep.txt
  CME213_STENCIL_G=$g python bench.py --workload stencil --extras 0 --no-cpu --no-e2e --steps 20 --warmup 3 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms %.4f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))" >> gpurun_out/r2_stencil_sweep.txt
done
python bench.py --workload c1 --extras 0 --no-cpu --no-e2e --steps 20 --warmup 3 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('c1 ms/pass %.5f frac %.3f copy floor ms %.5f' % (d['ms_per_step']/8, d['roofline']['frac'], d['config']['device_copy_of_the_same_bytes_ms_per_launch']))" >> gpurun_out/r2_stencil_sweep.txt
