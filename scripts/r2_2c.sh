#!/bin/bash
timeout 75 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 scripts/ddp_multistep_check.py 2>&1 | grep -E "ddp_multistep_check|Error|error|Traceback" | head -12
