#!/bin/bash
# e2e loader geometry sweep: prefetch group x groups ahead
for ga in "4 1" "4 2" "4 3" "8 1" "8 2" "2 3" "16 1"; do
  set -- $ga
  python bench.py --steps 4 --warmup 3 --no-cpu --group $1 --ahead $2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('group $1 ahead $2: value %.1fM e2e %.1fM strict_e2e %.1fM strict_dev %.1fM roofline %.3f' % (d['value']/1e6, d['e2e']['value']/1e6, d['e2e']['strict_alternation']/1e6, d['strict_alternation']['value']/1e6, d['roofline']['frac']))"
done
