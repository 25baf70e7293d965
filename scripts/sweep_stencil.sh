# stencil kernel shapes (tuning build: python 2012-04_stanford_cme213_b200/build.py --tuning)
# CME213_STENCIL_CFG = "lines per thread (0 = one-vector kernel), CTA threads, max threads along x, min CTAs per SM"
cd /root/repo 2>/dev/null || cd "$GRAFT_REPO_ROOT"
out=gpurun_out/${OUT:-r2_stencil_sweep2.txt}; : > $out
for cfg in ${CFGS:-"0,256,256,0" "1,256,256,0" "2,256,256,0" "2,128,128,0" "2,64,64,0" "2,256,128,0" "2,256,256,5" "2,256,256,6" "2,128,128,5" "1,256,256,8" "1,128,128,0" "1,128,128,8"}; do
  echo "== CME213_STENCIL_CFG=$cfg" >> $out
  CME213_STENCIL_CFG=$cfg python bench.py --workload stencil --extras 0 --no-cpu --no-e2e --steps 20 --warmup 3 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms %.4f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))" >> $out
done
