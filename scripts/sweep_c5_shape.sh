#!/bin/bash
# Launch-shape sweep of the lean CSR kernel on config C5 (512^3 fp64) and C1 (tuning build; GPU box).
# usage: bash scripts/sweep_c5_shape.sh > gpurun_out/c5_shape.txt
cd "$(dirname "$0")/.."
python 2012-04_stanford_cme213_b200/build.py --tuning > /dev/null || exit 1
for cfg in "" 64,512,0 64,1024,0 128,1024,0 128,1024,1 128,1536,0 128,2048,0 128,2048,1 128,1024,2 128,2048,2 256,2048,0 256,4096,0 512,4096,0; do
  echo "== c5 CME213_CSR_CFG=$cfg"
  CME213_CSR_CFG=$cfg python bench.py --workload c5 --csr-kernel 3 --steps 20 --warmup 3 --extras 0 --no-cpu --no-e2e 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms %.4f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))"
done
for cfg in "" 128,1024,0 256,2048,0 256,4096,0 512,4096,0 512,2048,0 1024,4096,0; do
  echo "== c1 CME213_CSR_CFG=$cfg"
  CME213_CSR_CFG=$cfg python bench.py --workload c1 --csr-kernel 3 --steps 50 --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms/pass %.5f frac %.3f' % (d['ms_per_step']/8, d['roofline']['frac']))"
done
python 2012-04_stanford_cme213_b200/build.py > /dev/null
