#!/bin/bash
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
MDT_NCCL_GRAPH=0 timeout 120 $TR --master-port 29551 scripts/nccl_h4096_check.py 2>&1 | grep -E "nccl_h4096_check|Error|error" | head -8
MDT_NCCL_GRAPH=0 timeout 200 python -m pytest tests/test_gpu_multi.py -m gpu -q --timeout=150 -p no:cacheprovider -k "two_groups or (four and nccl)" 2>&1 | tail -4
