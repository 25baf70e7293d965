#!/bin/bash
# Tile-shape sweep of the blocked bulk scan on the segmented fp32 scan (C2b) and the u32 scan (tuning build; GPU box).
cd "$(dirname "$0")/.."
python 2012-04_stanford_cme213_b200/build.py --tuning > /dev/null || exit 1
for cfg in "" 640,5,-4,2 768,4,-4,2 512,6,-4,2 704,4,-4,2 640,4,-5,3 576,5,-4,2 640,6,-3,1 512,5,-5,3 832,4,-4,2 1024,3,-4,2; do
  for w in segscan; do
    echo -n "$w CME213_SCAN_CFG=$cfg  "
    CME213_SCAN_CFG=$cfg python bench.py --workload $w --extras 0 --steps 30 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms %.4f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))"
  done
done
python 2012-04_stanford_cme213_b200/build.py > /dev/null
