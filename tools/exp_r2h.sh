# single GPU: cooperative node fetch A/B, then the per-config results of BASELINE.md section 5 (round 2)
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
for f in "-DRT3_NODE_COOP=1" "-DRT3_NODE_COOP=1 -DRT3_TRACE_MIN_BLOCKS=8"; do echo "== $f"; build "$f"; python -m pytest tests/test_gpu_parity.py -x -q -k "trace_closest" 2>&1 | tail -1; bench; done
build ""
echo "== run_configs"; python tools/run_configs.py cfg1 cfg3 cfg3_regen cfg4 cfg5 > gpurun_out/r2h_configs.jsonl 2> gpurun_out/r2h_configs.err; tail -3 gpurun_out/r2h_configs.err
echo "== run_next_rows"; python tools/run_next_rows.py > gpurun_out/r2h_next_rows.jsonl 2> gpurun_out/r2h_next_rows.err; tail -3 gpurun_out/r2h_next_rows.err
echo "== build times"; for m in "" rotations fast; do echo "-- $m"; python tools/build_times.py $m 2>&1 | tail -2; done; for m in "" rotations fast; do echo "-- grid 8 $m"; python tools/build_times.py 8 $m 2>&1 | tail -1; done
