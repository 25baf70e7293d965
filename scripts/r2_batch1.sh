cd /root/repo 2>/dev/null || cd "$GRAFT_REPO_ROOT"
timeout 500 python -m pytest tests -m gpu -x -q --timeout 200 2>&1 | tail -25 > gpurun_out/r2_pytest3.log
timeout 400 python scripts/sweep_csr.py > gpurun_out/r2_csr_sweep.csv 2> gpurun_out/r2_csr_sweep.err
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r2_bench_n1_a.json 2> gpurun_out/r2_bench_n1_a.err
timeout 500 bash scripts/sweep_c5_shape.sh > gpurun_out/r2_c5_shape.txt 2>&1
timeout 900 python scripts/ref_gpu_baseline.py > gpurun_out/r2_reference_gpu_programs.md 2> gpurun_out/r2_reference_gpu_programs.err
