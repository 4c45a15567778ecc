#!/bin/bash
# Throughput of both arms (this framework / stock PyTorch DDP+NCCL) for the given GPU counts, all three presets.
#   usage: tools/scaling_run.sh "2 4" gpurun_out/scaling.jsonl
# One JSON line per run is appended to the output file (bench.py's line, or the baseline harness' line).
set -u
NS=${1:-"1"}
OUT=${2:-gpurun_out/scaling.jsonl}
mkdir -p "$(dirname "$OUT")"
: > "$OUT"
PORT=29600
for N in $NS; do
  for P in baseline reference wide; do
    PORT=$((PORT + 1))
    if [ "$N" = "1" ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT"; fi
    timeout 300 $L bench.py --gpus "$N" --preset "$P" 2>>"$OUT.err" | grep '^{' >> "$OUT"
    PORT=$((PORT + 1))
    if [ "$N" = "1" ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT"; fi
    timeout 300 $L baseline/torch_ddp_baseline.py --preset "$P" --steps 100 2>>"$OUT.err" | grep '^{' >> "$OUT"
    PORT=$((PORT + 1))
    if [ "$N" = "1" ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT"; fi
    timeout 300 $L baseline/torch_ddp_baseline.py --preset "$P" --steps 100 --amp 2>>"$OUT.err" | grep '^{' >> "$OUT"
  done
done
python - "$OUT" <<'EOF'
import json, sys
for l in open(sys.argv[1]):
    d = json.loads(l)
    if d.get("impl") == "torch_ddp_baseline":
        print(f"ddp   N={d['n_gpus']} {d['preset']:9s} amp={d['amp_bf16']!s:5s} {d['images_per_sec']:12.0f} img/s  {d['ms_per_step']*1e3:8.1f} us/step")
    else:
        print(f"ours  N={d['n_gpus']} {d['config']['model'].split('(')[1].split(' ')[0]:9s}           {d['value']:12.0f} img/s  {d['ms_per_step']*1e3:8.1f} us/step  e2e {d['e2e']['value']:12.0f}")
EOF
