python -m pytest tests/test_gpu_kernels.py tests/test_gpu_engine.py -q -m gpu -x 2>&1 | tail -2
for rep in 1 2 3; do for p in 0 1; do
  echo "rep=$rep prefetch=$p: $(DTM_EPI_PREFETCH=$p python bench.py --no-presets --no-baseline --no-e2e --steps 400 --warmup 30 2>&1 | tail -1 | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(round(d["ms_per_step"]*1000,1))')"
done; done
