N=${1:-8}
cd /root/repo 2>/dev/null || cd "$GRAFT_REPO_ROOT"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for g in 1 0; do
timeout 300 $TR --master-port 2952$g bench.py --gpus $N --steps 40 --warmup 6 --graph $g --no-e2e --no-parity 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('N=$N graph=$g ms %.4f value %.0f kernel_ms %.4f per-rank uncoupled' % (d['ms_per_step'], d['value'], d['roofline']['kernel_ms']), ['%.4f' % v for v in (d['roofline'].get('kernel_ms_per_rank_uncoupled') or [])])"
done > gpurun_out/r2_multi_b_n$N.txt 2>&1
