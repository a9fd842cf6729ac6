# usage: bash tools/tp_scale.sh "2 4 8" [7b]   (inside gpurun --gpus N): collective alone (graph replay) + config-4 verify step per TP degree
for n in ${1:-2 4 8}; do
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n"
  timeout 120 $TR tools/tp_run.py ar 2>&1 | grep -E "TP_AR" | head -1
  timeout 300 $TR bench.py --config 4 --gpus $n --steps 20 --warmup 5 2>&1 | tail -1 | tee gpurun_out/r2_config4_tp$n.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('TP', d['n_gpus'], 'verify_us', round(d['value'],1), 'launches', d['gpu_launches'], 'err', d['tp_error_word'])"
  if [ -n "$2" ]; then timeout 300 $TR tools/tp_run.py bench 7b 20 2>&1 | grep TP_BENCH | cut -c1-110; fi
done
