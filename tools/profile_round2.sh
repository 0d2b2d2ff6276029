# Source-level ncu captures for round 2 (run AFTER the same command has exited 0 without ncu):
#   gpurun --timeout 900 -- 'bash tools/profile_round2.sh > gpurun_out/profile_round2.log 2>&1'
# then here:  python profiles/summarize.py source gpurun_out/r2_gbuffer.ncu-rep k_gbuffer   (and k_ref_shade, k_trace_dyn)
python -c "import __graft_entry__ as g; g.build()"
python bench.py --steps 3 --warmup 3 --no-extras > /dev/null 2>&1 || { echo "bench failed without ncu"; exit 1; }
for k in k_gbuffer k_ref_shade k_trace_dyn; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 6 -c 3 -o gpurun_out/r2_$k -f python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/ncu_$k.log 2>&1
  python profiles/summarize.py raw gpurun_out/r2_$k.ncu-rep $k > gpurun_out/r2_${k}_summary.txt 2>&1
  python profiles/summarize.py source gpurun_out/r2_$k.ncu-rep $k > gpurun_out/r2_${k}_source.txt 2>&1
done
