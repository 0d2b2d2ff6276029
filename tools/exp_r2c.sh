# GPU call of round 2: the strip / regeneration rewrite of the reference pass.  Full GPU suite, then RT3_STRIPS A/B of the bench.
line() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],3), 'ms; trace ms/launch', round(d['roofline']['kernel_ms_per_launch'],3), 'e2e', round(d['e2e']['value']), 'launches', d['gpu_launches'], 'parity', (d.get('parity') or {}).get('bit_identical'), (d.get('parity') or {}).get('host_frame_bit_identical'))
    elif 'Error' in l or 'error' in l or 'PARITY' in l: print(l.strip()[:300])
"; }
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for s in 1 2 3 4; do echo "== RT3_STRIPS=$s"; RT3_STRIPS=$s python bench.py --steps 40 --warmup 5 --no-extras 2>&1 | line; done
echo "== default, full line"; python bench.py --steps 40 --warmup 5 2>&1 | tee gpurun_out/r2c_bench.json | line
