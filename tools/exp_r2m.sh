python -c "import __graft_entry__ as g; g.build()"
t0=$(date +%s); python bench.py > gpurun_out/r2m_bench.json 2> gpurun_out/r2m_bench.err; tail -2 gpurun_out/r2m_bench.err; echo "bench wall $(( $(date +%s) - t0 )) s"
t0=$(date +%s); python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2m_bench_ref.json 2> gpurun_out/r2m_ref.err; tail -1 gpurun_out/r2m_ref.err; echo "reference arm wall $(( $(date +%s) - t0 )) s"
python -c "
import json
for f in ('gpurun_out/r2m_bench.json','gpurun_out/r2m_bench_ref.json'):
    for l in open(f):
        if l.startswith('{'):
            d=json.loads(l); print(f, round(d['value'],1), d['unit'], 'e2e', round(d['e2e']['value'],1), 'parity', (d.get('parity') or {}).get('bit_identical'), 'launches', d.get('gpu_launches'))
"
