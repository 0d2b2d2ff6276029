# ncu captures of the binary with the hit-byte-first node test (run AFTER the same command has exited 0 without ncu).
# gpurun --timeout 1500 -- 'bash tools/profile_r2w.sh'
python -c "import __graft_entry__ as g; g.build()"
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py > gpurun_out/r2w_bench.json 2> gpurun_out/r2w_bench.err || { echo "bench failed without ncu"; tail -5 gpurun_out/r2w_bench.err; exit 1; }
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2w_bench_ref.json 2> gpurun_out/r2w_ref.err; tail -1 gpurun_out/r2w_ref.err
python bench.py --steps 2 --warmup 3 --no-extras > /dev/null 2>&1 || { echo "short bench failed"; exit 1; }
# launch list: per-launch durations of the whole command (cold cache, serialised: compare SHARES)
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2w_launches.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r2w_ncu_launches.log 2>&1
python profiles/summarize.py launches gpurun_out/r2w_launches.csv > gpurun_out/r2w_launches_summary.txt 2>&1
for k in k_trace_dyn k_gbuffer k_ref_shade; do
  RT3_STRIPS=1 ncu --set full --clock-control none --import-source on -k regex:$k -s 6 -c 3 -o gpurun_out/r2w_$k -f python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/ncu_$k.log 2>&1
  python profiles/summarize.py raw gpurun_out/r2w_$k.ncu-rep $k > gpurun_out/r2w_${k}_summary.txt 2>&1
done
RAYS=$(python -c "
import json
for l in open('gpurun_out/r2w_bench.json'):
    if l.startswith('{'):
        d=json.loads(l); print(d['rays_per_step']-1920*1080)")
python profiles/summarize.py counters gpurun_out/r2w_k_trace_dyn.ncu-rep k_trace_dyn gpurun_out/r2w_trace_counters.json "the 3 traversal launches (bounces 1-3) of one 1080p frame of the bench workload with RT3_STRIPS = 1 (whole-frame queues, the launches bench.py times in its second pass), binary of profiles/r2w_bench.json (RT3_NODE_V2); rays = mean secondary rays per frame of that run" $RAYS > /dev/null 2>&1
python -c "
import json
for f in ('gpurun_out/r2w_bench.json','gpurun_out/r2w_bench_ref.json'):
    for l in open(f):
        if l.startswith('{'):
            d=json.loads(l); print(f, round(d['value'],1), d['unit'], 'e2e', round(d['e2e']['value'],1), 'parity', (d.get('parity') or {}).get('bit_identical'), 'cpu', d['cpu_baseline']['value'], d['cpu_baseline']['cores'], 'build', d['config'].get('build_ms'))
"
cat gpurun_out/r2w_launches_summary.txt | head -12
grep -E "time_duration|issue_active|thread_inst_executed_per|inst_executed.sum|warps_active|pipe_alu.avg|long_scoreboard|math_pipe" gpurun_out/r2w_k_trace_dyn_summary.txt
rm -f gpurun_out/r2w_k_gbuffer.ncu-rep gpurun_out/r2w_k_ref_shade.ncu-rep
