cd /root/repo 2>/dev/null || cd "$GRAFT_REPO_ROOT"
timeout 600 python -m pytest tests -m gpu -q --timeout 200 2>&1 | tail -40 > gpurun_out/r2_pytest4.log
for w in fp_iterate fp; do timeout 300 python bench.py --workload $w --extras 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['config']['workload'][:60], 'ms %.4f GB/s %.1f frac %.3f' % (d['ms_per_step'], d['value'], d['roofline']['frac']))"; done > gpurun_out/r2_fp.txt 2>&1
for k in 3 7; do timeout 300 python bench.py --workload c1 --csr-kernel $k --steps 50 --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('c1 kernel $k ms/pass %.5f frac %.3f' % (d['ms_per_step']/8, d['roofline']['frac']))"; done > gpurun_out/r2_pipe.txt 2>&1
for k in 3 7; do timeout 300 python bench.py --workload c5 --csr-kernel $k --extras 0 --no-cpu --no-e2e 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('c5 kernel $k ms %.4f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))"; done >> gpurun_out/r2_pipe.txt 2>&1
timeout 400 python scripts/sweep_csr.py > gpurun_out/r2_csr_sweep2.csv 2> gpurun_out/r2_csr_sweep2.err
timeout 900 python scripts/ref_gpu_baseline.py > gpurun_out/r2_reference_gpu_programs2.md 2> gpurun_out/r2_reference_gpu_programs2.err
