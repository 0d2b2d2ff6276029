# single GPU, final binary of round 2 (hit-byte-first node test + fast reciprocal): full GPU suite, BASELINE configs 1 / 3 / 4 / 5, next rows, build times, default bench + CPU arm
python -c "import __graft_entry__ as g; g.build()"
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/run_configs.py cfg1 cfg3 cfg3_regen cfg4 cfg5 > gpurun_out/r2z_configs.jsonl 2> gpurun_out/r2z_configs.err; tail -2 gpurun_out/r2z_configs.err
python tools/run_next_rows.py > gpurun_out/r2z_next_rows.jsonl 2> gpurun_out/r2z_next_rows.err; tail -2 gpurun_out/r2z_next_rows.err
(echo "-- 263 k"; python tools/build_times.py | tail -2; echo "-- 263 k rotations"; python tools/build_times.py rotations | tail -1; echo "-- 263 k fast"; python tools/build_times.py fast | tail -1; echo "-- 16.8 M"; python tools/build_times.py 8 | tail -1; echo "-- 16.8 M fast"; python tools/build_times.py 8 fast | tail -1) > gpurun_out/r2z_build_times.log 2>&1
python bench.py > gpurun_out/r2z_bench.json 2> gpurun_out/r2z_bench.err; tail -2 gpurun_out/r2z_bench.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2z_bench_ref.json 2> gpurun_out/r2z_ref.err; tail -1 gpurun_out/r2z_ref.err
python -c "
import json
for f in ('gpurun_out/r2z_bench.json','gpurun_out/r2z_bench_ref.json'):
    for l in open(f):
        if l.startswith('{'):
            d=json.loads(l); print(f, round(d['value'],1), d['unit'], 'e2e', round(d['e2e']['value'],1), 'parity', (d.get('parity') or {}).get('bit_identical'), 'cpu', d['cpu_baseline']['value'], d['cpu_baseline']['cores'], 'build', d['config'].get('build_ms'))
"
cut -c1-300 gpurun_out/r2z_configs.jsonl; cat gpurun_out/r2z_build_times.log
