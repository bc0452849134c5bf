#!/bin/bash
# tuning sweep: prints us/step and the calibrated per-kernel times for a few knob settings
run() { echo "== $*"; env "$@" timeout 200 python bench.py --steps 3000 --warmup 20 --no-cpu-baseline --no-modes 2>/dev/null | python -c "
import json,sys; j=json.loads(sys.stdin.read()); print(round(j['ms_per_step']*1e3,2), 'us/step |', ' '.join(f\"{k['kernel']}={k['avg_us']}\" for k in j['kernels']))"; }
run A=0
run MDT_BN_FC4=32
run MDT_BN_FC4=32 MDT_MAX_SPLITS=3
run MDT_MAX_SPLITS=3
run MDT_MAX_SPLITS=2
run MDT_BN_H=64
run MDT_BN_SKINNY=32
run MDT_NO_SPLITK=1
