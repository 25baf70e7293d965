cd /root/repo 2>/dev/null || cd "$GRAFT_REPO_ROOT"
timeout 600 python -m pytest tests -m gpu -q --timeout 200 2>&1 | tail -8 > gpurun_out/r2_pytest8.log
timeout 900 python bench.py > gpurun_out/r2_bench_n1_c.json 2> gpurun_out/r2_bench_n1_c.err
timeout 900 python scripts/ref_gpu_baseline.py > gpurun_out/r2_reference_gpu_programs6.md 2> gpurun_out/r2_reference_gpu_programs6.err
