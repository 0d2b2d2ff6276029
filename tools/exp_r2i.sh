# single GPU: regeneration as a dense kernel (k_ref_regen), present pass, full suite
python -c "import __graft_entry__ as g; g.build()"
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/run_configs.py cfg3_regen > gpurun_out/r2i_regen.jsonl 2> gpurun_out/r2i_regen.err; tail -2 gpurun_out/r2i_regen.err; cat gpurun_out/r2i_regen.jsonl
python bench.py --steps 40 --warmup 5 > gpurun_out/r2i_bench.json 2> gpurun_out/r2i_bench.err; tail -2 gpurun_out/r2i_bench.err
python -c "
import json
for l in open('gpurun_out/r2i_bench.json'):
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), 'Mrays/s e2e', round(d['e2e']['value']), 'strong_base', d['strong_base'])
"
