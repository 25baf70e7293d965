cd /root/repo 2>/dev/null || cd "$GRAFT_REPO_ROOT"
timeout 600 python -m pytest tests -m gpu -q --timeout 200 2>&1 | tail -40 > gpurun_out/r2_pytest5.log
timeout 400 python scripts/sweep_csr.py > gpurun_out/r2_csr_sweep3.csv 2> gpurun_out/r2_csr_sweep3.err
for w in hyb hyb_window csr_powerlaw csr_window fp_iterate pagerank; do timeout 300 python bench.py --workload $w --extras 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['config']['workload'][:90], 'ms %.4f GB/s %.1f frac %.3f' % (d['ms_per_step'], d['value'], d['roofline']['frac']))"; done > gpurun_out/r2_c4.txt 2>&1
timeout 300 python bench.py --workload hyb --hyb-tail-kernel 3 --extras 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['config']['workload'][:90], 'ms %.4f' % d['ms_per_step'])" >> gpurun_out/r2_c4.txt 2>&1
timeout 300 python bench.py --workload hyb_window --hyb-tail-kernel 3 --extras 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['config']['workload'][:90], 'ms %.4f' % d['ms_per_step'])" >> gpurun_out/r2_c4.txt 2>&1
for K in 3 5 7 9 12 16; do timeout 300 python bench.py --workload hyb_window --hyb-k $K --extras 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('banded K=$K ms %.4f' % d['ms_per_step'])"; timeout 300 python bench.py --workload hyb --hyb-k $K --extras 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('uniform K=$K ms %.4f' % d['ms_per_step'])"; done >> gpurun_out/r2_c4.txt 2>&1
python 2012-04_stanford_cme213_b200/build.py --tuning > /dev/null
for cfg in "" 640,3,-3,1 640,2,-4,2 768,2,-4,2 512,3,-4,2 1024,2,-3,1 896,2,-3,1; do echo "fp_iterate CME213_SCAN_CFG=$cfg"; CME213_SCAN_CFG=$cfg timeout 300 python bench.py --workload fp_iterate --extras 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms %.4f' % d['ms_per_step'])"; done > gpurun_out/r2_fp_shapes.txt 2>&1
python 2012-04_stanford_cme213_b200/build.py > /dev/null
timeout 900 python scripts/ref_gpu_baseline.py > gpurun_out/r2_reference_gpu_programs3.md 2> gpurun_out/r2_reference_gpu_programs3.err
