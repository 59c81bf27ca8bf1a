#!/bin/bash
# GPU box: A/B of library variants muscle_b200/lib/exp/libmuscle_b200_<n>.so given as arguments
mkdir -p gpurun_out
cp muscle_b200/lib/libmuscle_b200.so /tmp/lib_orig.so
for n in "$@"; do
  cp muscle_b200/lib/exp/libmuscle_b200_$n.so muscle_b200/lib/libmuscle_b200.so
  echo "== variant $n"
  timeout 300 python scripts/pair_ab.py config2 config3 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: print(l.rstrip()); continue
    print('%s: rowsum_sym %.1f us  rowsum_ordered %.1f  grad %.1f  step %.1f' % (d['workload'], d.get('rowsum_sym_us', 0), d['rowsum_ordered_us'], d['grad_us'], d['step_us']))"
done
cp /tmp/lib_orig.so muscle_b200/lib/libmuscle_b200.so
