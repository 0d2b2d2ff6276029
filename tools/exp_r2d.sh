# 2-GPU call: bench.py --gpus 2 (strong scaling of configs[2]) with both delivery paths.   gpurun --gpus 2 -- 'bash tools/exp_r2d.sh'
N=${1:-2}
python -c "import __graft_entry__ as g; g.build()"
for mode in p2p a2a; do
  echo "== --gpus $N --exchange $mode"
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 3 --warmup 3 --exchange $mode > gpurun_out/r2d_n${N}_$mode.log 2>&1
  echo "rc=$?"; grep '^{' gpurun_out/r2d_n${N}_$mode.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],2), 'ms/step; e2e', round(d['e2e']['value']), 'delivery', d['config']['delivery'], d['config']['delivery_note'], 'verified', d['gathered_equals_single_gpu_image'], 'single', round(d['single_gpu_same_workload']['value']), 'weak', round(d['weak']['value']), round(d['weak']['ms_per_step'],3), 'balance', round(d['config']['rays_max_over_mean'],4))
"
  grep -v '^{' gpurun_out/r2d_n${N}_$mode.log | grep -i "error\|Traceback\|assert\|PARITY" | head -5
done
