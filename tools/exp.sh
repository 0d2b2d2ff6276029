# usage: tools/exp.sh "<nvcc extra flags>" ... : rebuild with each flag set on the GPU box and print bench numbers (experiments only)
for flags in "$@"; do
  echo "== RT3_NVCC_EXTRA=$flags"
  RT3_NVCC_EXTRA="$flags" python -c "from raytracer3_b200 import build as b; b.build(force=True, verbose=True)" 2>&1 | grep -A3 "k_trace_dynILb0ELi28ELb0" | grep "Used\|spill"
  python -m pytest tests/test_gpu_parity.py -x -q -k "trace_closest or reference_pass" 2>&1 | tail -1
  python bench.py --steps 30 --warmup 5 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],3), 'ms; trace ms/launch', round(d['roofline']['kernel_ms_per_launch'],3), 'incoh', round(d['incoherent']['closest_mrays_s']), round(d['incoherent']['any_mrays_s']), 'e2e', round(d['e2e']['value']))
"
done
