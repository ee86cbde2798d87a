#!/bin/bash
# The BASELINE configs at scale on one 8xB200 box (gpurun --gpus 8): config 3 (bench.py, fixed 20 000-genome job),
# config 4 (100 000 phage genomes, sparse list), config 5 (500 x 5 Mb) at N = 8 and, side by side on disjoint GPUs,
# at N = 4 / 2 / 1; then the drop-in binary end to end on 8 GPUs and on 1.  Outputs: gpurun_out/r02_*.
cd "$(dirname "$0")/.."
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nvidia-smi --query-gpu=index,name,memory.total --format=csv,noheader > $O/r02_8gpu_box.txt; nproc >> $O/r02_8gpu_box.txt; free -g | head -2 >> $O/r02_8gpu_box.txt
date +%T
timeout 600 $TR --nproc-per-node 8 --master-port 29601 bench.py --gpus 8 --config 3 --steps 2 --warmup 1 > $O/r02_config3_n8.jsonl 2> $O/r02_config3_n8.err
echo "config3 rc=$?"; date +%T
timeout 120 python bench.py --impl reference --gpus 8 --config 3 --steps 2 --warmup 1 > $O/r02_config3_n8_reference_arm.jsonl 2> $O/r02_config3_ref.err
echo "config3 ref rc=$?"; date +%T
timeout 600 $TR --nproc-per-node 8 --master-port 29602 tools/configs45.py --config 4 > $O/r02_config4_n8.jsonl 2> $O/r02_config4_n8.err
echo "config4 rc=$?"; date +%T
timeout 600 $TR --nproc-per-node 8 --master-port 29603 tools/configs45.py --config 5 --steps 1 --warmup 1 > $O/r02_config5_n8.jsonl 2> $O/r02_config5_n8.err
echo "config5 n8 rc=$?"; date +%T
( CUDA_VISIBLE_DEVICES=0,1,2,3 timeout 900 $TR --nproc-per-node 4 --master-port 29604 tools/configs45.py --config 5 --steps 1 --warmup 0 --no-e2e > $O/r02_config5_n4.jsonl 2> $O/r02_config5_n4.err; echo "config5 n4 rc=$?" ) &
( CUDA_VISIBLE_DEVICES=4,5 timeout 900 $TR --nproc-per-node 2 --master-port 29605 tools/configs45.py --config 5 --steps 1 --warmup 0 --no-e2e > $O/r02_config5_n2.jsonl 2> $O/r02_config5_n2.err; echo "config5 n2 rc=$?" ) &
( CUDA_VISIBLE_DEVICES=6 timeout 900 python tools/configs45.py --config 5 --steps 1 --warmup 0 --no-e2e > $O/r02_config5_n1.jsonl 2> $O/r02_config5_n1.err; echo "config5 n1 rc=$?" ) &
# meanwhile GPU 7: the drop-in binary on one GPU (2 000-genome FASTA, reference CLI on the first 300 genomes)
( CUDA_VISIBLE_DEVICES=7 timeout 600 tools/product_e2e.sh 2000 0 300 > $O/r02_product_e2e_1gpu.txt 2>&1; echo "product 1gpu rc=$?" ) &
wait
date +%T
timeout 300 tools/product_e2e.sh 2000 0,1,2,3,4,5,6,7 > $O/r02_product_e2e_8gpu.txt 2>&1
echo "product 8gpu rc=$?"; date +%T
tail -2 $O/r02_config3_n8.err $O/r02_config4_n8.err $O/r02_config5_n8.err $O/r02_config5_n1.err | cut -c1-300
