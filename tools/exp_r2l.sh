# single GPU: launch fixed cost / tail measured directly, a per-rank frame of the 8-GPU run on one GPU with its kernel list, steady-state DRAM traffic
python -c "import __graft_entry__ as g; g.build()"
python tools/tail_stats.py sweep > gpurun_out/r2l_tail_sweep.jsonl 2> gpurun_out/r2l_tail.err; tail -2 gpurun_out/r2l_tail.err; cat gpurun_out/r2l_tail_sweep.jsonl
python tools/tail_stats.py rankframe > gpurun_out/r2l_rankframe.jsonl 2>> gpurun_out/r2l_tail.err; cat gpurun_out/r2l_rankframe.jsonl
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2l_rankframe_launches.csv python tools/tail_stats.py rankframe > /dev/null 2>&1
python profiles/summarize.py launches gpurun_out/r2l_rankframe_launches.csv
# DRAM traffic of the frame's kernels WITHOUT ncu's cache flush between replays (steady state: scene, BVH and shading records stay in L2)
python bench.py --steps 2 --warmup 3 --no-extras > /dev/null 2>&1
RT3_STRIPS=1 ncu --cache-control none --clock-control none --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct -k regex:'k_trace_dyn|k_ref_shade|k_ref_init|k_ref_finish|k_gbuffer' -s 12 -c 9 --csv --log-file gpurun_out/r2l_frame_dram_nocacheflush.csv python bench.py --steps 3 --warmup 3 --no-extras > /dev/null 2>&1
tail -40 gpurun_out/r2l_frame_dram_nocacheflush.csv | cut -c1-260
