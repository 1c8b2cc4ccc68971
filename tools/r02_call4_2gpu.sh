#!/bin/bash
# Round 2, GPU call 4 (TWO B200s): the 2-rank NCCL test of the three exchanges (gathered tiles, tiles stored into rank 0's
# frame over NVLink peer memory, reduce-scatter of sample ranges), tools/p2p_check.py, and 2-GPU bench lines with phases.
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
nvidia-smi topo -m > $O/r02d_topo.txt 2>&1
( timeout 600 python -m pytest tests/test_gpu_properties.py -m gpu -q -k "two_gpu or exported or resolve_pack or resolve_range" > $O/r02d_pytest_2gpu.txt 2>&1; echo "pytest exit $?" >> $O/r02d_pytest_2gpu.txt )
tail -5 $O/r02d_pytest_2gpu.txt
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/p2p_check.py > $O/r02d_p2p_check.log 2>&1
grep -E "^\{" $O/r02d_p2p_check.log | cut -c1-300
export NS="2" STEPS=20
tools/scale_run.sh $O/r02d_c1_w1_tiles.jsonl --world 1 --pp 16 --width 1280 --height 720
tools/scale_run.sh $O/r02d_c1_w1_tiles_p2p.jsonl --world 1 --pp 16 --width 1280 --height 720 --shard tiles_p2p
tools/scale_run.sh $O/r02d_c3_w3_samples.jsonl --world 3 --pp 32 --width 1280 --height 720 --shard samples
tools/scale_run.sh $O/r02d_w5_1080p_tiles.jsonl
tools/scale_run.sh $O/r02d_w5_1080p_tiles_p2p.jsonl --shard tiles_p2p
