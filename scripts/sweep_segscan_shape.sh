#!/bin/bash
# Tile-shape sweep of the blocked bulk scan on the segmented fp32 scan (C2b) and the u32 scan (tuning build; GPU box).
cd "$(dirname "$0")/.."
python 2012-04_stanford_cme213_b200/build.py --tuning > /dev/null || exit 1
for cfg in 384,7,-5,3 320,7,-6,4 256,7,-7,5 384,5,-6,4 448,5,-5,3 384,7,-4,2; do
  CME213_SCAN_CFG=$cfg python bench.py --workload scan --extras 0 --steps 30 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('scan u32 $cfg ms %.4f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))"
  for w in segscan; do
    echo -n "$w CME213_SCAN_CFG=$cfg  "
    CME213_SCAN_CFG=$cfg python bench.py --workload $w --extras 0 --steps 30 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms %.4f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))"
  done
done
for cfg in 576,3,-4,2 384,3,-6,4 320,3,-7,5; do
  echo -n "fp_iterate CME213_SCAN_CFG=$cfg  "
  CME213_SCAN_CFG=$cfg python bench.py --workload fp_iterate --extras 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms %.4f' % d['ms_per_step'])"
done
python 2012-04_stanford_cme213_b200/build.py > /dev/null
