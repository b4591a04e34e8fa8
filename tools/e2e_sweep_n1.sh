#!/bin/bash
# end-to-end arm at N = 1 under different worker / sub-batch / chunk settings (bench.py env knobs)
run(){ tag=$1; shift; env "$@" python bench.py --steps 6 --warmup 3 --no-configs --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); e=d['e2e']; print('$tag', round(e['value'],2), 'GB/s', round(e['ms_per_step'],1), 'ms', 'prov', round(e['provider_only']['value'],1), 'cons', round(e['consumer_only']['value'],1), 'frac', round(e['frac_of_pcie'],3))"; }
for a in "$@"; do run $a $(echo $a | tr ',' ' '); done
