# treelet leaf-count threshold (RT3_TREELET_MIN_LEAVES): build time and frame rate on the GPU
line() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],3), 'ms; trace ms/launch', round(d['roofline']['kernel_ms_per_launch'],4), 'build ms', round(d['config']['build_ms'],2), 'nodes', d['config']['wide_nodes'])
    elif 'Error' in l or 'error' in l or 'PARITY' in l: print(l.strip()[:300])
"; }
for m in 0 7 16 32; do
  echo "== RT3_TREELET_MIN_LEAVES=$m"
  RT3_NVCC_EXTRA="-DRT3_TREELET_MIN_LEAVES=$m" python -c "from raytracer3_b200 import build as b; b.build(force=True)"
  python tools/build_times.py 2>&1 | tail -1 | cut -c1-260
  python bench.py --steps 40 --warmup 5 --no-extras 2>&1 | line
done
echo "== RT3_TREELET_MIN_LEAVES=7, forced treelets on the 16.8 M-triangle grid"
RT3_NVCC_EXTRA="-DRT3_TREELET_MIN_LEAVES=7 -DRT3_TREELET_AUTO_MAX=0xFFFFFFFFu" python -c "from raytracer3_b200 import build as b; b.build(force=True)"
python tools/build_times.py 8 2>&1 | tail -1 | cut -c1-260
python -c "from raytracer3_b200 import build as b; b.build(force=True)"
