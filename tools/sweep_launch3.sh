for cfg in "12 512 37" "16 320 37" "16 384 37" "14 512 37" "12 512 42" "16 256 37"; do
  set -- $cfg
  python bench.py --steps 600 --warmup 5 --quick --pipeline $1 --threads $2 --grid $3 2>/dev/null | python -c "
import sys,json
j=json.loads(sys.stdin.read()); print('cfg $cfg', 'value', round(j['value'],1), 'frac', round(j['roofline']['frac'],3), 'kms', round(j['roofline']['kernel_ms'],3), 'resident', j['roofline']['concurrent_launches'])"
done
