# single-GPU A/B: reciprocal ray direction by MUFU.RCP + Newton step (RT3_FAST_RCP, default) against three IEEE divisions per fetched ray; KEEP sweep on the new kernel
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
echo "== -DRT3_FAST_RCP=0"; build "-DRT3_FAST_RCP=0"; bench; bench
echo "== default build (RT3_FAST_RCP=1)"; build ""
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
bench; bench
for k in 24 32; do echo "== RT3_TRACE_KEEP=$k"; RT3_TRACE_KEEP=$k bench; done
python tools/run_configs.py cfg4 2>/dev/null | cut -c1-330 | tail -1
