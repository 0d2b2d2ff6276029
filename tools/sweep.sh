# usage: tools/sweep.sh : sweep of the traversal tunables (lanes kept busy before a warp refills, resident blocks per SM) on the GPU box
for cfg in "28 8" "24 8" "30 8" "32 8" "28 6" "28 10" "28 12" "26 10"; do set -- $cfg; echo -n "KEEP=$1 BLOCKS_PER_SM=$2: "; RT3_TRACE_KEEP=$1 RT3_TRACE_BLOCKS_PER_SM=$2 python bench.py --steps 40 --warmup 5 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value']), 'Mrays/s', round(d['ms_per_step'],3), 'ms; trace ms/launch', round(d['roofline']['kernel_ms_per_launch'],4), 'incoh', round(d['incoherent']['closest_mrays_s']), round(d['incoherent']['any_mrays_s']))
"; done
