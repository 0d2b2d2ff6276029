# ncu captures of the current binary (run AFTER the same command has exited 0 without ncu).  TAG names the output files.
# gpurun --timeout 1500 -- 'TAG=r2ae bash tools/profile_tag.sh'
TAG=${TAG:-r2ae}
python -c "import __graft_entry__ as g; g.build()"
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || { echo "bench failed without ncu"; tail -5 gpurun_out/${TAG}_bench.err; exit 1; }
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_ref.err; tail -1 gpurun_out/${TAG}_ref.err
python bench.py --steps 2 --warmup 3 --no-extras > /dev/null 2>&1 || { echo "short bench failed"; exit 1; }
# launch list: per-launch durations of the whole command (cold cache, serialised: compare SHARES)
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/${TAG}_ncu_launches.log 2>&1
python profiles/summarize.py launches gpurun_out/${TAG}_launches.csv > gpurun_out/${TAG}_launches_summary.txt 2>&1
for k in k_trace_dyn k_gbuffer k_ref_shade; do
  RT3_STRIPS=1 ncu --set full --clock-control none --import-source on -k regex:$k -s 6 -c 3 -o gpurun_out/${TAG}_$k -f python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/ncu_$k.log 2>&1
  python profiles/summarize.py raw gpurun_out/${TAG}_$k.ncu-rep $k > gpurun_out/${TAG}_${k}_summary.txt 2>&1
done
RAYS=$(python -c "
import json
for l in open('gpurun_out/${TAG}_bench.json'):
    if l.startswith('{'):
        d=json.loads(l); print(d['rays_per_step']-1920*1080)")
python profiles/summarize.py counters gpurun_out/${TAG}_k_trace_dyn.ncu-rep k_trace_dyn gpurun_out/${TAG}_trace_counters.json "the 3 traversal launches (bounces 1-3) of one 1080p frame of the bench workload with RT3_STRIPS = 1 (whole-frame queues, the launches bench.py times in its second pass), binary of profiles/${TAG}_bench.json (RT3_NODE_V2 + RT3_FAST_RCP: the final binary of round 2); rays = mean secondary rays per frame of that run" $RAYS > /dev/null 2>&1
python -c "
import json
for f in ('gpurun_out/${TAG}_bench.json','gpurun_out/${TAG}_bench_ref.json'):
    for l in open(f):
        if l.startswith('{'):
            d=json.loads(l); print(f, round(d['value'],1), d['unit'], 'e2e', round(d['e2e']['value'],1), 'parity', (d.get('parity') or {}).get('bit_identical'), 'cpu', d['cpu_baseline']['value'], d['cpu_baseline']['cores'], 'build', d['config'].get('build_ms'))
"
cat gpurun_out/${TAG}_launches_summary.txt | head -12
grep -E "time_duration|issue_active|thread_inst_executed_per|inst_executed.sum|warps_active|pipe_alu.avg|long_scoreboard|math_pipe" gpurun_out/${TAG}_k_trace_dyn_summary.txt
rm -f gpurun_out/${TAG}_k_gbuffer.ncu-rep gpurun_out/${TAG}_k_ref_shade.ncu-rep
