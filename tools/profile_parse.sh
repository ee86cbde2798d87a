#!/bin/bash
# usage: tools/profile_parse.sh <tag>   -- ncu full capture of k_parse_pairs on a 300-genome run + region breakdown
set -e
TAG=$1
cd /root/repo
/usr/local/graft/bin/gpurun --timeout 900 -- "python bench.py --genomes 300 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_parse -s 1 -c 1 -o gpurun_out/${TAG}_parse python bench.py --genomes 300 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu2.log 2>&1; tail -1 gpurun_out/ncu2.log" 2>&1 | tail -2
ncu -i gpurun_out/${TAG}_parse.ncu-rep --page source --csv > gpurun_out/${TAG}_source.csv 2>/dev/null
ncu -i gpurun_out/${TAG}_parse.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
mkdir -p /tmp/sass && cd /tmp/sass && rm -f *.cubin && cuobjdump -xelf all /root/repo/lz-ani_b200/liblzani_b200.so >/dev/null && nvdisasm -g -c kernels.sm_100a.cubin > k.sass
cd /root/repo
python tools/ncu_regions.py gpurun_out/${TAG}_source.csv /tmp/sass/k.sass PrmDefault 89700
python - <<PY
import csv
rows=list(csv.reader(open('gpurun_out/${TAG}_raw.csv')))
d=dict(zip(rows[0],rows[2]))
for k in ['gpu__time_duration.sum','smsp__inst_executed.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio','smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio','smsp__average_warps_issue_stalled_wait_per_issue_active.ratio','smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio','smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio','l1tex__throughput.avg.pct_of_peak_sustained_elapsed','launch__registers_per_thread','smsp__thread_inst_executed_per_inst_executed.ratio','dram__bytes_read.sum','dram__bytes_write.sum']:
    print(k, d.get(k))
PY
