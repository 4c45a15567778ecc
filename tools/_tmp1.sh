for rep in 1 2; do for t in 256 384 512; do
  echo "rep=$rep update_threads=$t: $(DTM_UPDATE_THREADS=$t python bench.py --no-presets --no-baseline --no-e2e --steps 400 --warmup 30 2>&1 | tail -1 | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(round(d["ms_per_step"]*1000,1))')"
done; done
B="python bench.py --no-graph --no-e2e --no-baseline --no-presets --steps 2 --warmup 3"
$B > gpurun_out/ncu_plain.log 2>&1 && ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:dtm:: -s 60 -c 24 -f -o gpurun_out/step_kernels_r2 $B > gpurun_out/ncu_run.log 2>&1
tail -3 gpurun_out/ncu_run.log; ls -la gpurun_out/*.ncu-rep
