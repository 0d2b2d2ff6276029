"""Summaries of ncu outputs: `python profiles/summarize.py launches <csv>`, `... raw <ncu-rep> [kernel-substring]` or `... source <ncu-rep> [kernel-substring]`."""
import collections
import csv
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__inst_executed.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'lts__t_bytes.sum', 'lts__t_sectors.sum', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct', 'l1tex__t_bytes.sum', 'smsp__sass_average_branch_targets_threads_uniform.pct',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__inst_executed_pipe_alu.sum', 'smsp__inst_executed_pipe_fma.sum', 'smsp__inst_executed_pipe_lsu.sum', 'smsp__inst_executed_pipe_xu.sum',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed_op_local_ld.sum', 'smsp__inst_executed_op_local_st.sum', 'local_load_requests', 'smsp__cycles_active.avg']


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    ki, vi = hdr.index('Kernel Name'), hdr.index('Metric Value')
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        try:
            v = float(r[vi].replace(',', ''))
        except ValueError:
            continue
        n = r[ki].split('(')[0][-48:]
        agg[n][0] += 1
        agg[n][1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"# {path}: {sum(v[0] for v in agg.values())} launches, {tot / 1e6:.3f} ms total (ncu: cold cache, serialised; compare SHARES)")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:16]:
        print(f"{v[1] / 1e6:9.3f} ms {v[0]:4d} launches {100 * v[1] / tot:5.1f}%  avg {v[1] / v[0] / 1e3:8.1f} us  {k}")


def raw(path, sub=None):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    data = [r for r in rows[2:] if sub is None or sub in r[hdr.index('Kernel Name')]]
    print('# kernels:', [r[hdr.index('Kernel Name')][:40] for r in data])
    for k in KEYS:
        if k in hdr:
            i = hdr.index(k)
            print(f"{k:80s} {units[i]:14s}", [r[i] for r in data])


def source(path, sub=None, top=40):
    """`... source <ncu-rep> [kernel-substring]`: executed warp instructions per SOURCE LINE of the first matching kernel (needs -lineinfo and
    `--import-source on`), heaviest lines first — where a kernel spends its instructions, e.g. the attribute half of k_gbuffer."""
    hdr = rows = None
    for extra in (['--print-source', 'cuda'], []):                 # CUDA-C lines if the report carries the source, else SASS lines
        args = ['ncu', '-i', path, '--page', 'source', '--csv'] + extra + (['-k', 'regex:' + sub] if sub else [])
        out = subprocess.run(args, capture_output=True, text=True).stdout
        rows = [r for r in csv.reader(out.splitlines()) if len(r) > 3]
        hdr = next((r for r in rows if any(c.strip() == 'Source' for c in r) and any('Instructions Executed' in c for c in r)), None)
        if hdr is not None:
            break
    if hdr is None:
        print('no source page in', path, '(column names differ? look at: ncu -i', path, '--page source --csv | head)')
        return
    si = next(i for i, c in enumerate(hdr) if c.strip() == 'Source')
    ii = next((i for i, c in enumerate(hdr) if c.strip() == 'Instructions Executed'), None)
    if ii is None:
        ii = next(i for i, c in enumerate(hdr) if 'Instructions Executed' in c)
    agg = collections.defaultdict(float)
    for r in rows[rows.index(hdr) + 1:]:
        try:
            agg[r[si].strip()[:150]] += float(r[ii].replace(',', ''))
        except (ValueError, IndexError):
            continue
    tot = sum(agg.values()) or 1.0
    print(f"# {path} {sub or ''}: {tot / 1e6:.1f} M warp instructions")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:top]:
        print(f"{100 * v / tot:5.1f}%  {k}")


def counters(path, sub, out_json, note="", rays_total=None):
    """`... counters <ncu-rep> <kernel-substring> <out.json> [note]`: per-launch ncu counters of a kernel as JSON for bench.py's roofline objects
    (DRAM bytes, L2 bytes, warp instructions, active lanes; means over the captured launches of the kernel)."""
    import json
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    data = [r for r in rows[2:] if sub in r[hdr.index('Kernel Name')]]

    def col(name, scale_by_unit=True):
        if name not in hdr:
            return None
        i = hdr.index(name)
        mult = {'Mbyte': 1e6, 'Gbyte': 1e9, 'Kbyte': 1e3, 'byte': 1.0, 'us': 1e-6, 'ms': 1e-3, 'ns': 1e-9}.get(units[i], 1.0) if scale_by_unit else 1.0
        return [float(r[i].replace(',', '')) * mult for r in data]
    n = len(data)
    mean = lambda v: sum(v) / len(v) if v else None
    lts = col('lts__t_sectors.sum', False)                                   # 32-byte sectors at the L2 (all sources); lts__t_bytes is not in `--set full`
    dr, dw, lt, inst, lanes, dur = col('dram__bytes_read.sum'), col('dram__bytes_write.sum'), col('lts__t_bytes.sum') or ([32.0 * x for x in lts] if lts else None), col('smsp__inst_executed.sum'), \
        col('smsp__thread_inst_executed_per_inst_executed.ratio'), col('gpu__time_duration.sum')
    git = subprocess.run(['git', 'rev-parse', '--short', 'HEAD'], capture_output=True, text=True).stdout.strip()
    rec = {"source": f"{out_json.split('/')[-1]} <- ncu --set full capture {path.split('/')[-1]} of kernel '{sub}', {n} launches, summarised at commit {git}" + (f"; {note}" if note else ""),
           "launches": n, "dram_bytes_per_launch": mean([a + b for a, b in zip(dr, dw)]) if dr and dw else None,
           "dram_read_bytes": dr, "dram_write_bytes": dw, "l2_bytes_per_launch": mean(lt) if lt else None,
           "warp_inst_per_launch": mean(inst) if inst else None, "warp_inst": inst, "active_lanes_per_inst": mean(lanes) if lanes else None,
           "ncu_duration_s": dur,
           "rays_in_captured_launches": rays_total, "warp_inst_per_ray": (sum(inst) / rays_total) if (inst and rays_total) else None,
           "dram_bytes_per_ray": (sum(dr) + sum(dw)) / rays_total if (dr and dw and rays_total) else None,
           "l2_bytes_per_ray": sum(lt) / rays_total if (lt and rays_total) else None}
    with open(out_json, 'w') as f:
        json.dump(rec, f, indent=1)
    print(json.dumps(rec, indent=1))


if __name__ == '__main__':
    if sys.argv[1] == 'launches':
        launches(sys.argv[2])
    elif sys.argv[1] == 'counters':
        counters(sys.argv[2], sys.argv[3], sys.argv[4], sys.argv[5] if len(sys.argv) > 5 else "", float(sys.argv[6]) if len(sys.argv) > 6 else None)
    elif sys.argv[1] == 'source':
        source(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else None)
    else:
        raw(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else None)

