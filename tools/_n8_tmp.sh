T="timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$T --master-port 29511 tools/multi_gpu_check.py --sweep --out gpurun_out/multi_gpu_8_r2.json 2>&1 | grep -E "FAIL|RESULT|fused_allreduce" | cut -c1-260
for cfg in "0 auto" "64 auto" "32 auto" "0 0"; do set -- $cfg
  echo "overlap_ctas=$1 nvls=$2: $(DTM_OVERLAP_CTAS=$1 DTM_NVLS=$2 $T --master-port 29513 bench.py --gpus 8 --steps 300 --warmup 20 --no-baseline --no-presets --no-e2e 2>/dev/null | grep metric | python -c 'import sys,json; print(round(json.loads(sys.stdin.read())["ms_per_step"]*1000,1))')"
done
$T --master-port 29512 tools/timeline.py 2>&1 | grep -v Warn | tail -8
