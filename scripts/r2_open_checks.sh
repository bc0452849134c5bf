#!/bin/bash
# The GPU checks round 2 ended without (its GPU budget was spent): run on a box with >= 2 B200s.
#   1. multi-step graphs vs single steps on a DDP group, every exchange, both GEMM modes
#   2. the optim.SGD drop-in loop against plain torch-CUDA (tests/test_gpu_dropin.py, opt-in test)
#   3. the same comparison as 1. inside the bench, with its fallbacks (key "multistep_parity" of the JSON line)
set -u
cd "$(dirname "$0")/.."
N=${N:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port 29631 \
    scripts/ddp_multistep_check.py; echo "ddp_multistep_check rc=$?"
MDT_TEST_UNVALIDATED=1 timeout 600 python -m pytest tests/test_gpu_dropin.py -q -m gpu -k sgd; echo "sgd drop-in rc=$?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port 29632 \
    bench.py --gpus "$N" --steps 100 --warmup 10 --no-eager --no-modes | python -c "
import json, sys
line = [l for l in sys.stdin if l.startswith('{')][-1]
d = json.loads(line)
print(json.dumps({k: d.get(k) for k in ('multistep_parity', 'ddp_parity', 'layouts')}, indent=1))"
