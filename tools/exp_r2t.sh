# head check: full GPU suite, next rows (post-process with the division-free log2), default bench + CPU arm
python -c "import __graft_entry__ as g; g.build()"
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/run_next_rows.py > gpurun_out/r2t_next_rows.jsonl 2> gpurun_out/r2t_next_rows.err; tail -2 gpurun_out/r2t_next_rows.err; head -2 gpurun_out/r2t_next_rows.jsonl | cut -c1-260
python bench.py > gpurun_out/r2t_bench.json 2> gpurun_out/r2t_bench.err; tail -2 gpurun_out/r2t_bench.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2t_bench_ref.json 2> gpurun_out/r2t_ref.err; tail -1 gpurun_out/r2t_ref.err
python -c "
import json
for f in ('gpurun_out/r2t_bench.json','gpurun_out/r2t_bench_ref.json'):
    for l in open(f):
        if l.startswith('{'):
            d=json.loads(l); print(f, round(d['value'],1), d['unit'], 'e2e', round(d['e2e']['value'],1), 'parity', (d.get('parity') or {}).get('bit_identical'), 'cpu', d['cpu_baseline']['value'], d['cpu_baseline']['cores'], 'build', d['config'].get('build_ms'))
"
