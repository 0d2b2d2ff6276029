# warp-cooperative treelet restructuring: correctness (full suite), build stage times, bench
python -c "import __graft_entry__ as g; g.build()"
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for m in "" rotations; do echo "-- 263 k $m"; python tools/build_times.py $m 2>&1 | tail -2 | cut -c1-260; done
echo "-- 16.8 M (default = rotations above 4 M records)"; python tools/build_times.py 8 2>&1 | tail -1 | cut -c1-260
echo "-- 16.8 M treelets forced"; RT3_NVCC_EXTRA="-DRT3_TREELET_AUTO_MAX=0xFFFFFFFFu" python -c "from raytracer3_b200 import build as b; b.build(force=True)"; python tools/build_times.py 8 2>&1 | tail -1 | cut -c1-260
python -c "from raytracer3_b200 import build as b; b.build(force=True)"
python bench.py --steps 40 --warmup 5 --no-extras 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],3), 'ms; build ms', round(d['config']['build_ms'],2), 'nodes', d['config']['wide_nodes'])
    elif 'rror' in l: print(l.strip()[:300])
"
