cd /root/repo 2>/dev/null || cd "$GRAFT_REPO_ROOT"
timeout 400 python -m pytest tests/test_sharded_gpu.py tests/test_zz_no_alloc.py -m gpu -q --timeout 150 2>&1 | tail -60 > gpurun_out/r2_dbg2.log
timeout 200 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 150 2>&1 | tail -30 >> gpurun_out/r2_dbg2.log
timeout 300 2012-04_stanford_cme213_b200/bin/csr_iterate --gpus 2 --grid 256 --iters 10 --mode split --check >> gpurun_out/r2_dbg2.log 2>&1
timeout 300 python bench.py --workload pagerank_iterate --extras 0 --steps 5 > gpurun_out/r2_pr_iter.json 2>/dev/null
