#!/bin/bash
# GPU box: times every built experiment variant (scripts/build_variants.sh) on config 2 and config 3.
cd "$(dirname "$0")/.."
cp muscle_b200/lib/libmuscle_b200.so /tmp/lib_orig.so
for so in muscle_b200/lib/exp/libmuscle_b200_*.so; do
  cp $so muscle_b200/lib/libmuscle_b200.so
  c2=$(timeout 120 python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('%.4f ms grad %.1f us rowsum %.1f us' % (d['ms_per_step'], d['roofline']['us_per_launch'], d['roofline_rowsum']['us_per_launch']))")
  c3=$(timeout 120 python bench.py --workload config3 --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('%.4f ms grad %.1f us rowsum %.1f us' % (d['ms_per_step'], d['roofline']['us_per_launch'], d['roofline_rowsum']['us_per_launch']))")
  echo "$(basename $so) | config2: $c2 | config3: $c3"
done
cp /tmp/lib_orig.so muscle_b200/lib/libmuscle_b200.so
