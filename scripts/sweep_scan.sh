# usage: bash scripts/sweep_scan.sh   (on the GPU box) -- sweeps tile shapes of the u32 scan
for c in default 256,4,0 1024,8,0 512,4,2 512,4,3 512,6,2 512,8,2 512,8,3 1024,4,2 1024,4,3 1024,6,2 1024,8,1 1024,2,4 1024,3,4 1024,5,2 1024,7,1; do
  v=$(CME213_SCAN_CFG=$c timeout 40 python bench.py --workload scan --steps 20 --warmup 3 --extras 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['ms_per_step'],4), round(d['roofline']['frac'],3))")
  echo "cfg $c: $v"
done
for w in segscan fp hyb; do timeout 40 python bench.py --workload $w --steps 20 --warmup 3 --extras 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$w', round(d['value'],1), round(d['ms_per_step'],4), round(d['roofline']['frac'],3))"; done
