# single-GPU A/Bs: chunked ray fetch, resident blocks, 64-byte shading records, RT3_TRI_WAIT
line() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],3), 'ms; trace ms/launch (strips=1 pass)', round(d['roofline']['kernel_ms_per_launch'],4), 'e2e', round(d['e2e']['value']))
    elif 'Error' in l or 'error' in l or 'PARITY' in l: print(l.strip()[:300])
"; }
build() { RT3_NVCC_EXTRA="$1" python -c "from raytracer3_b200 import build as b; b.build(force=True)"; }
bench() { python bench.py --steps 40 --warmup 5 --no-extras 2>&1 | line; }
echo "== default build"; build ""; python -m pytest tests -m gpu -x -q 2>&1 | tail -2; bench
echo "== RT3_TRACE_BLOCKS_PER_SM=9"; RT3_TRACE_BLOCKS_PER_SM=9 bench
echo "== RT3_TRACE_BLOCKS_PER_SM=10"; RT3_TRACE_BLOCKS_PER_SM=10 bench
echo "== RT3_STRIPS=1"; RT3_STRIPS=1 bench
for f in "-DRT3_FETCH_CHUNK=0" "-DRT3_FETCH_CHUNK=32" "-DRT3_FETCH_CHUNK=128" "-DRT3_SHADE64=0" "-DRT3_SHADE_MIN_BLOCKS=4" "-DRT3_TRI_WAIT=8"; do echo "== $f"; build "$f"; bench; done
echo "== -DRT3_TRI_WAIT=8 BLOCKS_PER_SM=9"; RT3_TRACE_BLOCKS_PER_SM=9 bench
build ""
