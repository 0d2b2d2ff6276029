# NOTE: the -DRT3_NODE96 code was removed after this measurement; it is profiles/r2ai_node96_experiment.patch
# single-GPU A/B: 96-byte nodes fetched with three 256-bit loads (-DRT3_NODE96=1) against 80-byte nodes with five 128-bit loads
line() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],3), 'ms; trace ms/launch (strips=1 pass)', round(d['roofline']['kernel_ms_per_launch'],4), 'e2e', round(d['e2e']['value']), 'parity', (d.get('parity') or {}).get('bit_identical'))
    elif 'Error' in l or 'error' in l or 'PARITY' in l: print(l.strip()[:300])
"; }
build() { RT3_NVCC_EXTRA="$1" python -c "from raytracer3_b200 import build as b; b.build(force=True)"; }
bench() { python bench.py --steps 40 --warmup 5 --no-extras 2>&1 | line; }
echo "== default build"; build ""; bench; bench
echo "== -DRT3_NODE96=1"; build "-DRT3_NODE96=1"; bench; bench
python bench.py --steps 20 --warmup 5 2>&1 | line
python tools/run_configs.py cfg1 cfg4 2>/dev/null | cut -c1-330 | tail -2
build ""
