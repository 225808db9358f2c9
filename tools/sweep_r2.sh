#!/usr/bin/env bash
# pipeline depth x block size sweep at the driver's run shape (20 steps) and at 200 steps
for steps in 20 200; do
for cfg in "2 1024" "3 1024" "4 1024" "4 512" "6 512" "8 512" "12 512" "16 512" "6 1024"; do
  set -- $cfg
  v=$(python bench.py --steps $steps --warmup 5 --quick --pipeline $1 --threads $2 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['roofline']['kernel_ms'],3), round(d['roofline']['kernel_share_of_step'],2))")
  echo "steps $steps pipeline $1 threads $2 -> $v"
done; done
