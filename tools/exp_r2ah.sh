# single-GPU A/B: prefetch of the next child node at the end of a node step (-DRT3_PREFETCH_NEXT=1)
line() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],3), 'ms; trace ms/launch (strips=1 pass)', round(d['roofline']['kernel_ms_per_launch'],4), 'e2e', round(d['e2e']['value']))
    elif 'Error' in l or 'error' in l or 'PARITY' in l: print(l.strip()[:300])
"; }
build() { RT3_NVCC_EXTRA="$1" python -c "from raytracer3_b200 import build as b; b.build(force=True)"; }
bench() { python bench.py --steps 40 --warmup 5 --no-extras 2>&1 | line; }
echo "== default build"; build ""; bench; bench
echo "== -DRT3_PREFETCH_NEXT=1 (the flag and its code were removed after this measurement: two prefetch.global.L1 of the next child at the end of the node step)"; build "-DRT3_PREFETCH_NEXT=1"; bench; bench
python tools/run_configs.py cfg4 2>/dev/null | cut -c1-330 | tail -1
build ""
