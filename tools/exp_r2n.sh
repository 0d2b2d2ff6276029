python -c "import __graft_entry__ as g; g.build()"
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for i in 1 2; do python bench.py --steps 60 --warmup 5 > gpurun_out/r2n_bench_$i.json 2> gpurun_out/r2n_bench.err; tail -1 gpurun_out/r2n_bench.err | cut -c1-200
python -c "
import json
for l in open('gpurun_out/r2n_bench_$i.json'):
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'ms', round(d['e2e']['ms_per_step'],3), 'sync', round(d['e2e']['sync_ms_per_step'],3), 'parity', d['parity']['bit_identical'], d['parity']['host_frame_bit_identical'])
"; done
