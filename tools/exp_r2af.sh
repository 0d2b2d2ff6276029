# single GPU: work-index -> pixel map by multiply-high (pass_types.h fastdiv) — full GPU suite, bench, next rows (post-process)
python -c "import __graft_entry__ as g; g.build()"
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for i in 1 2; do python bench.py --steps 40 --warmup 5 --no-extras 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],3), 'ms; e2e', round(d['e2e']['value']))
"; done
python tools/run_next_rows.py 2>/dev/null | cut -c1-260 | head -3
python tools/run_configs.py cfg1 2>/dev/null | cut -c1-200
