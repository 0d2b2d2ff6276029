python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for cfg in "28 1" "24 1"; do set -- $cfg; echo "KEEP=$1 TRIS=$2"; RT3_TRACE_KEEP=$1 RT3_TRACE_TRIS=$2 python bench.py --steps 30 --warmup 5 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],3), 'ms; trace ms/launch', round(d['roofline']['kernel_ms_per_launch'],3), 'incoh', round(d['incoherent']['closest_mrays_s']), round(d['incoherent']['any_mrays_s']), 'e2e', round(d['e2e']['value']))
"; done
