#!/bin/bash
# BASELINE configs 4 and 5 at scale on one 8xB200 box (gpurun --gpus 8), after the hashed matrix scan / PrmLong changes:
# config 5 at N = 8, then N = 4 / 2 / 1 side by side on disjoint GPUs (config 4 at N = 1 on the eighth), then config 4
# at N = 8.  Outputs: gpurun_out/r02z_*  (r02x_* = the same script run before the chunk growth and the adaptive hand-over).
cd "$(dirname "$0")/.."
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
date +%T
timeout 600 $TR --nproc-per-node 8 --master-port 29603 tools/configs45.py --config 5 --steps 1 --warmup 1 > $O/r02z_config5_n8.jsonl 2> $O/r02z_config5_n8.err
echo "config5 n8 rc=$?"; date +%T
( CUDA_VISIBLE_DEVICES=0,1,2,3 timeout 900 $TR --nproc-per-node 4 --master-port 29604 tools/configs45.py --config 5 --steps 1 --warmup 0 --no-e2e > $O/r02z_config5_n4.jsonl 2> $O/r02z_config5_n4.err; echo "config5 n4 rc=$?" ) &
( CUDA_VISIBLE_DEVICES=4,5 timeout 900 $TR --nproc-per-node 2 --master-port 29605 tools/configs45.py --config 5 --steps 1 --warmup 0 --no-e2e > $O/r02z_config5_n2.jsonl 2> $O/r02z_config5_n2.err; echo "config5 n2 rc=$?" ) &
( CUDA_VISIBLE_DEVICES=6 timeout 900 python tools/configs45.py --config 5 --steps 1 --warmup 0 --no-e2e > $O/r02z_config5_n1.jsonl 2> $O/r02z_config5_n1.err; echo "config5 n1 rc=$?" ) &
( CUDA_VISIBLE_DEVICES=7 timeout 900 python tools/configs45.py --config 4 > $O/r02z_config4_n1.jsonl 2> $O/r02z_config4_n1.err; echo "config4 n1 rc=$?" ) &
wait
date +%T
timeout 600 $TR --nproc-per-node 8 --master-port 29602 tools/configs45.py --config 4 > $O/r02z_config4_n8.jsonl 2> $O/r02z_config4_n8.err
echo "config4 n8 rc=$?"; date +%T
for f in $O/r02z_config*.jsonl; do echo $f; grep '^{' $f | cut -c1-60,230-420; done
tail -n 2 $O/r02z_config5_n8.err | cut -c1-300
