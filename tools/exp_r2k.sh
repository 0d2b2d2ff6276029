line() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],3), 'ms; trace ms/launch (strips=1 pass)', round(d['roofline']['kernel_ms_per_launch'],4), 'e2e', round(d['e2e']['value']))
    elif 'Error' in l or 'error' in l or 'PARITY' in l: print(l.strip()[:300])
"; }
build() { RT3_NVCC_EXTRA="$1" python -c "from raytracer3_b200 import build as b; b.build(force=True)"; }
bench() { python bench.py --steps 40 --warmup 5 --no-extras 2>&1 | line; }
echo "== default build"; build ""; bench
echo "== -DRT3_TRACE_THREADS=64 (16 CTAs/SM)"; build "-DRT3_TRACE_THREADS=64"; python -m pytest tests/test_gpu_parity.py -x -q -k "trace_closest" 2>&1 | tail -1; bench
echo "== -DRT3_TRACE_THREADS=64 RT3_STRIPS=3"; RT3_STRIPS=3 bench
echo "== -DRT3_TRACE_THREADS=256 (4 CTAs/SM)"; build "-DRT3_TRACE_THREADS=256"; bench
build ""
