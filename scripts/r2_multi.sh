# multi-GPU batch: bash scripts/r2_multi.sh N   (run under gpurun --gpus N)
N=${1:-2}
cd /root/repo 2>/dev/null || cd "$GRAFT_REPO_ROOT"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 python -m pytest tests/test_sharded_gpu.py tests/test_zz_no_alloc.py -m gpu -q --timeout 150 2>&1 | grep -E "Error|error|passed|failed|assert" | head -20 > gpurun_out/r2_sharded_n$N.log
timeout 500 $TR --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err
echo "bench rc=$?" >> gpurun_out/r2_sharded_n$N.log
timeout 400 $TR --master-port 29513 scripts/verify_sharded.py --grid 256 --iters 6 > gpurun_out/r2_verify_sharded_n$N.log 2>&1
echo "verify rc=$?" >> gpurun_out/r2_sharded_n$N.log
timeout 300 $TR --master-port 29514 scripts/diag_h2d.py > gpurun_out/r2_diag_h2d_n$N.log 2>&1
timeout 300 2012-04_stanford_cme213_b200/bin/csr_iterate --gpus $N --grid 384 --iters 20 --check > gpurun_out/r2_csr_iterate_n$N.log 2>&1
for m in gather split; do timeout 300 2012-04_stanford_cme213_b200/bin/csr_iterate --gpus $N --grid 256 --iters 10 --mode $m --check >> gpurun_out/r2_csr_iterate_n$N.log 2>&1; done
timeout 400 $TR --master-port 29515 bench.py --gpus $N --steps 20 --warmup 5 --exchange fused --no-e2e > gpurun_out/r2_bench_n${N}_fused.json 2> gpurun_out/r2_bench_n${N}_fused.err
