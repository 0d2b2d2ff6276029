# N-GPU call: BASELINE configs[4] as written (1080p frame of the 8x8 grid over N GPUs), both build modes.   gpurun --gpus N -- 'bash tools/exp_r2z_cfg5.sh N'
N=${1:-8}
python -c "import __graft_entry__ as g; g.build()"
for MODE in "" "--two-level"; do
  TAG=$(echo "$MODE" | tr -d " -")
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2959$N bench.py --gpus $N --workload cfg5 $MODE --steps 40 --warmup 5 > gpurun_out/r2z_cfg5_n${N}${TAG}.log 2>&1
  echo "rc=$?"; grep '^{' gpurun_out/r2z_cfg5_n${N}${TAG}.log | tee gpurun_out/r2z_cfg5_n${N}${TAG}.json | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],3), 'ms/step; e2e', round(d['e2e']['value']), 'verified', d.get('gathered_equals_single_gpu_image'), 'single', d.get('single_gpu_same_workload') and round(d['single_gpu_same_workload']['ms_per_step'],3), 'build ms', round(d['config']['build_ms'],1))
"
  grep -v '^{' gpurun_out/r2z_cfg5_n${N}${TAG}.log | grep -i "error\|Traceback\|assert" | head -3
done
