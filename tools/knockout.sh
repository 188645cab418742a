#!/bin/bash
# In-graph marginal cost of each kernel of the decode step: step time with that launch left out (B200_DEBUG_SKIP, model.cu).
# Usage (GPU box): bash tools/knockout.sh > gpurun_out/knockout.txt
run() { B200_DEBUG_SKIP=$1 timeout 300 python bench.py --steps 20 --warmup 5 --skip-cpu-baseline --no-engine 2>/dev/null | tail -n 1 > /tmp/ko.json; python -c "
import json,sys
d=json.load(open('/tmp/ko.json')); print('skip %4s %-28s ms_per_step %.3f  sm_mhz %s' % (sys.argv[1], sys.argv[2], d['ms_per_step'], d['clocks']['sm_mhz']))" $1 "$2"; }
run 0 "(nothing)"
run 1 ln_shift_mix6
run 2 "rkv+lora-down group"
run 4 "lora-up group"
run 8 "fused wkv"
run 16 "o-proj"
run 32 ln_shift_mix1
run 64 "ffn key"
run 128 "ffn value"
run 256 "ln_out+head"
run 0 "(nothing, again)"
