cd /root/repo 2>/dev/null || cd "$GRAFT_REPO_ROOT"
timeout 600 python -m pytest tests -m gpu -q --timeout 200 2>&1 | tail -30 > gpurun_out/r2_pytest6.log
bash scripts/r2_ncu.sh
timeout 900 python scripts/ref_gpu_baseline.py > gpurun_out/r2_reference_gpu_programs4.md 2> gpurun_out/r2_reference_gpu_programs4.err
