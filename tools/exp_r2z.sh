# N-GPU call: bench.py --gpus N, default settings, full line kept.   gpurun --gpus N -- 'bash tools/exp_r2z.sh N'
N=${1:-2}; shift; EXTRA="$*"; TAG=$(echo "$EXTRA" | tr -d " -")
python -c "import __graft_entry__ as g; g.build()"
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2958$N bench.py --gpus $N --steps ${STEPS:-5} --warmup 3 $EXTRA > gpurun_out/r2z_n${N}${TAG}.log 2>&1
echo "rc=$?"; grep '^{' gpurun_out/r2z_n${N}${TAG}.log | tee gpurun_out/r2z_n${N}${TAG}.json | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],2), 'ms/step; e2e', round(d['e2e']['value']), 'delivery', d['config']['delivery'], d['config']['delivery_note'], 'verified', d['gathered_equals_single_gpu_image'], 'single', round(d['single_gpu_same_workload']['value']), 'weak', d['weak'] and round(d['weak']['value']), 'inframe', d['inframe_64spp'] and round(d['inframe_64spp']['ms_per_step'],2), 'build ms', round(d['config']['build_ms'],1))
"
grep -v '^{' gpurun_out/r2z_n${N}${TAG}.log | grep -i "error\|Traceback\|assert\|PARITY" | head -5
