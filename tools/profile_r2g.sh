# ncu captures of the round-2 binary (run AFTER the same command has exited 0 without ncu).  gpurun --timeout 1200 -- 'bash tools/profile_r2g.sh'
python -c "import __graft_entry__ as g; g.build()"
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --steps 40 --warmup 5 > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err || { echo "bench failed without ncu"; tail -5 gpurun_out/r2g_bench.err; exit 1; }
python bench.py --steps 2 --warmup 3 --no-extras > /dev/null 2>&1 || { echo "short bench failed"; exit 1; }
# launch list: per-launch durations of the whole command (cold cache, serialised: compare SHARES)
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2g_launches.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r2g_ncu_launches.log 2>&1
python profiles/summarize.py launches gpurun_out/r2g_launches.csv > gpurun_out/r2g_launches_summary.txt 2>&1
for k in k_gbuffer k_ref_shade k_trace_dyn k_ref_init; do
  RT3_STRIPS=1 ncu --set full --clock-control none --import-source on -k regex:$k -s 6 -c 3 -o gpurun_out/r2g_$k -f python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/ncu_$k.log 2>&1
  python profiles/summarize.py raw gpurun_out/r2g_$k.ncu-rep $k > gpurun_out/r2g_${k}_summary.txt 2>&1
done
