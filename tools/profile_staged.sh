#!/bin/bash
# usage: tools/profile_staged.sh <tag> [families members]  -- ncu full capture of the walk and record kernels of the
# staged parse on a small set in one chunk (default 15 families x 20 = the bench mix) + per-region breakdown
set -e
TAG=$1; NF=${2:-15}; MEM=${3:-20}
cd /root/repo
/usr/local/graft/bin/gpurun --timeout 900 -- "export LZANI_STAGED_WORDS=200000000; python tools/run_set.py $NF $MEM 2 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_walk|k_hit_spec|k_hit_probe' -c 3 -o gpurun_out/${TAG}_staged python tools/run_set.py $NF $MEM > gpurun_out/ncu2.log 2>&1; cat gpurun_out/plain.log" 2>&1 | tail -5
mkdir -p /tmp/sass && cd /tmp/sass && rm -f *.cubin && cuobjdump -xelf all /root/repo/lz-ani_b200/liblzani_b200.so >/dev/null && nvdisasm -g -c kernels.sm_100a.cubin > k.sass
cd /root/repo
ncu -i gpurun_out/${TAG}_staged.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
NP=$((NF*MEM*(NF*MEM-1)))
for K in k_hit_probe k_hit_spec k_walk_pairs; do
  ncu -i gpurun_out/${TAG}_staged.ncu-rep --page source --csv -k regex:$K > /tmp/${TAG}_${K}_all.csv 2>/dev/null
  n=$(grep -n '^"Kernel Name"' /tmp/${TAG}_${K}_all.csv | sed -n 2p | cut -d: -f1)
  if [ -n "$n" ]; then head -n $((n-1)) /tmp/${TAG}_${K}_all.csv > gpurun_out/${TAG}_${K}_source.csv; else cp /tmp/${TAG}_${K}_all.csv gpurun_out/${TAG}_${K}_source.csv; fi
  echo "=== $K"
  python tools/ncu_regions.py gpurun_out/${TAG}_${K}_source.csv /tmp/sass/k.sass "${K}INS_10PrmDefault" $NP
done
