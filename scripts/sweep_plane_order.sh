#!/bin/bash
# CTA order of the stream kernel on config C5 (512^3, fp64): strip width (row blocks) x planes per chunk (tuning build; GPU box).
cd "$(dirname "$0")/.."
python 2012-04_stanford_cme213_b200/build.py --tuning > /dev/null || exit 1
for cfg in "-1,0" "64,8" "32,8" "128,8" "256,8" "64,4" "64,16" "64,32" "16,8" "2048,8" "64,512" "512,4"; do
  echo -n "c5 CME213_PLANE_ORDER=$cfg  "
  CME213_PLANE_ORDER=$cfg python bench.py --workload c5 --steps 20 --warmup 3 --extras 0 --no-cpu --no-e2e 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms %.4f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))"
done
python 2012-04_stanford_cme213_b200/build.py > /dev/null
