for cfg in "8 512 37" "10 512 29" "6 512 49" "8 512 35" "9 512 32" "12 384 37"; do
  set -- $cfg
  python bench.py --steps 400 --warmup 5 --quick --pipeline $1 --threads $2 --grid $3 2>/dev/null | python -c "
import sys,json
j=json.loads(sys.stdin.read()); print('cfg $cfg', 'value', round(j['value'],1), 'frac', round(j['roofline']['frac'],3), 'kms', round(j['roofline']['kernel_ms'],3))"
done
