#!/bin/bash
# Repeats the driver-shaped device-resident leg (bench.py --steps 20 --warmup 5 --quick) to see its run-to-run spread, with and
# without the delayed first NVML query of the clock sampler.  Prints value and the number of clock samples per run.
for d in 0.0012 0; do
  echo "delay=$d"
  for i in 1 2 3 4 5 6; do
    DMS_CLOCK_SAMPLER_DELAY_S=$d python bench.py --steps 20 --warmup 5 --quick 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(j['value']), j['clocks'].get('samples'))"
  done
done
