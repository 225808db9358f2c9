for cfg in "16 256 32" "18 256 29" "16 256 34" "20 256 26" "24 256 22" "12 384 42"; do
  set -- $cfg
  python bench.py --steps 600 --warmup 5 --quick --pipeline $1 --threads $2 --grid $3 2>/dev/null | python -c "
import sys,json
j=json.loads(sys.stdin.read()); print('cfg $cfg', 'value', round(j['value'],1), 'frac', round(j['roofline']['frac'],3), 'kms', round(j['roofline']['kernel_ms'],3), 'resident', j['roofline']['concurrent_launches'])"
done
