T="timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$T --master-port 29511 tools/multi_gpu_check.py --out gpurun_out/multi_gpu_8_r2_final_checks.json 2>&1 | grep -E "FAIL|RESULT" | cut -c1-600
$T --master-port 29513 bench.py --gpus 8 --steps 200 --warmup 20 > gpurun_out/r2_bench_n8.json 2> gpurun_out/r2_bench_n8.err
grep metric gpurun_out/r2_bench_n8.json | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"], d["vs_baseline"], d["e2e"]["ms_per_step"], d["fused_exchange"], d["presets"])'
$T --master-port 29512 tools/timeline.py 2>&1 | grep -v Warn | tail -15 > gpurun_out/timeline_8gpu_baseline_r2.txt; tail -6 gpurun_out/timeline_8gpu_baseline_r2.txt
