# First GPU call of round 2 (everything here was written after round 1's GPU minutes ran out and has only run on the CPU emulations):
#   gpurun --timeout 1500 -- 'bash tools/exp_round2.sh > gpurun_out/exp_round2.log 2>&1'
# 1. the env-gated GPU tests of the opt-in features, 2. RT3_BUILD_TREELET A/B of the whole bench, 3. the traversal-kernel experiments.
line() { python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],3), 'ms; trace ms/launch', round(d['roofline']['kernel_ms_per_launch'],3), 'incoh', round(d['incoherent']['closest_mrays_s']), round(d['incoherent']['any_mrays_s']), 'build ms', round(d['config']['build_ms'],2), 'nodes', d['config']['wide_nodes'])
"; }
python -c "import __graft_entry__ as g; g.build()"
echo "== unverified GPU tests"; RT3_TEST_UNVERIFIED=1 python -m pytest tests/test_gpu_unverified.py -m gpu -x -q 2>&1 | tail -3
echo "== default";              python bench.py --steps 30 --warmup 5 2>&1 | line
echo "== RT3_FORCE_TREELET=1";  RT3_FORCE_TREELET=1 python bench.py --steps 30 --warmup 5 2>&1 | line
bash tools/exp.sh "-DRT3_TRI_WAIT=8" "-DRT3_TRI_WAIT=6" "-DRT3_TRI_WAIT=10" "-DRT3_TRI_BALANCE=1 -DRT3_TRI_WAIT=16 -DRT3_TRACE_MIN_BLOCKS=8" "-DRT3_SHADE_MIN_BLOCKS=5" "-DRT3_SHADE_MIN_BLOCKS=6"
echo "== RT3_TRI_WAIT=8 + treelet"; RT3_NVCC_EXTRA="-DRT3_TRI_WAIT=8" python -c "from raytracer3_b200 import build as b; b.build(force=True)"; RT3_FORCE_TREELET=1 python bench.py --steps 30 --warmup 5 2>&1 | line
