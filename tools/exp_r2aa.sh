# single-GPU A/B: second record step only when >= T lanes still hold a record (-DRT3_TRI_REP2_MIN=T)
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
echo "== default build"; build ""; bench; bench
for t in 2 3 4 6; do echo "== -DRT3_TRI_REP2_MIN=$t"; build "-DRT3_TRI_REP2_MIN=$t"; bench; done
echo "== -DRT3_TRI_REP2_MIN=3: tests + cfg4"; build "-DRT3_TRI_REP2_MIN=3"; python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -1; python tools/run_configs.py cfg4 2>/dev/null | cut -c1-330 | tail -1
build ""
