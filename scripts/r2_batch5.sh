cd /root/repo 2>/dev/null || cd "$GRAFT_REPO_ROOT"
timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 150 2>&1 | tail -5 > gpurun_out/r2_pytest7.log
timeout 600 python -m pytest tests -m gpu -q --timeout 200 2>&1 | tail -8 >> gpurun_out/r2_pytest7.log
timeout 900 python scripts/ref_gpu_baseline.py > gpurun_out/r2_reference_gpu_programs5.md 2> gpurun_out/r2_reference_gpu_programs5.err
timeout 900 python bench.py > gpurun_out/r2_bench_n1_b.json 2> gpurun_out/r2_bench_n1_b.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1
