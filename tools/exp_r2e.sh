# 8-GPU call: bench.py --gpus 8 (strong scaling of configs[2]).   gpurun --gpus 8 -- 'bash tools/exp_r2e.sh'
N=${1:-8}
python -c "import __graft_entry__ as g; g.build()"
run() {
  echo "== --gpus $N $* (RT3_STRIPS=${RT3_STRIPS:-default})"
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N --steps 5 --warmup 3 "$@" > gpurun_out/r2e_tmp.log 2>&1
  echo "rc=$?"; grep '^{' gpurun_out/r2e_tmp.log | tee -a gpurun_out/r2e_n${N}.jsonl | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],2), 'ms/step; e2e', round(d['e2e']['value']), 'delivery', d['config']['delivery'], d['config']['delivery_note'], 'verified', d['gathered_equals_single_gpu_image'], 'single', round(d['single_gpu_same_workload']['value']), 'weak', round(d['weak']['value']), round(d['weak']['ms_per_step'],3), 'balance', round(d['config']['rays_max_over_mean'],4), 'trace ms/launch', round(d['roofline']['kernel_ms_per_launch'],4), 'share', round(d['roofline']['kernel_share_of_step'],3))
"
  grep -v '^{' gpurun_out/r2e_tmp.log | grep -i "error\|Traceback\|assert\|PARITY" | head -5
}
run
RT3_STRIPS=1 run
run --tile 32
nvidia-smi topo -m 2>/dev/null | head -12
