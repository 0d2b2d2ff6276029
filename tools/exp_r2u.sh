# NOTE: the -DRT3_HALF_CVT code was removed after this measurement; it is profiles/r2u_half_cvt_experiment.patch
# single-GPU A/Bs (round 2, session 13): resident traversal CTAs x strips (room for the shade kernels of the other strip beside a
# persistent traversal grid), and -DRT3_HALF_CVT=1 (byte -> float conversions of the node test on the FMA pipe through FHFMA)
line() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],3), 'ms; trace ms/launch (strips=1 pass)', round(d['roofline']['kernel_ms_per_launch'],4), 'e2e', round(d['e2e']['value']), 'parity', (d.get('parity') or {}).get('bit_identical'))
    elif 'Error' in l or 'error' in l or 'PARITY' in l: print(l.strip()[:300])
"; }
build() { RT3_NVCC_EXTRA="$1" python -c "from raytracer3_b200 import build as b; b.build(force=True)"; }
bench() { python bench.py --steps 40 --warmup 5 --no-extras 2>&1 | line; }
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
echo "== default build"; build ""; bench; bench
for b in 7 6 5; do for s in 2 3; do echo "== RT3_TRACE_BLOCKS_PER_SM=$b RT3_STRIPS=$s"; RT3_TRACE_BLOCKS_PER_SM=$b RT3_STRIPS=$s bench; done; done
echo "== -DRT3_HALF_CVT=1 (72 registers) RT3_TRACE_BLOCKS_PER_SM=7"; build "-DRT3_HALF_CVT=1"; RT3_TRACE_BLOCKS_PER_SM=7 bench
echo "== -DRT3_HALF_CVT=1 -DRT3_TRACE_MIN_BLOCKS=8 (64 registers)"; build "-DRT3_HALF_CVT=1 -DRT3_TRACE_MIN_BLOCKS=8"
python -m pytest tests -m gpu -x -q 2>&1 | tail -2; bench; bench
python tools/run_configs.py cfg1 cfg4 2>/dev/null | cut -c1-400 | tail -4
echo "== default build again"; build ""; bench
