# single-GPU A/B: RT3_NODE_V2 (hit byte first: valid24 node word, permutation table in shared memory) against round 1's node format
line() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],3), 'ms; trace ms/launch (strips=1 pass)', round(d['roofline']['kernel_ms_per_launch'],4), 'e2e', round(d['e2e']['value']))
    elif 'Error' in l or 'error' in l or 'PARITY' in l: print(l.strip()[:300])
"; }
build() { RT3_NVCC_EXTRA="$1" python -c "from raytracer3_b200 import build as b; b.build(force=True)"; }
bench() { python bench.py --steps 40 --warmup 5 --no-extras 2>&1 | line; }
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
echo "== -DRT3_NODE_V2=0 (round 1's node format)"; build "-DRT3_NODE_V2=0"; bench; bench
python tools/run_configs.py cfg1 cfg4 2>/dev/null | cut -c1-330 | tail -2
echo "== default build (RT3_NODE_V2=1)"; build ""
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
bench; bench
python tools/run_configs.py cfg1 cfg4 cfg5 2>/dev/null | cut -c1-420 | tail -3
python tools/run_next_rows.py 2>/dev/null | cut -c1-300 | head -6
