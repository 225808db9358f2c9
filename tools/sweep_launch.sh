for cfg in "6 0 0" "6 512 49" "12 512 24" "8 512 37" "12 512 0" "16 256 37" "12 256 49" "24 256 24"; do
  set -- $cfg
  python bench.py --steps 240 --warmup 3 --quick --pipeline $1 --threads $2 --grid $3 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    try: j=json.loads(l)
    except: continue
    print('cfg $cfg', 'value', round(j['value'],1), 'frac', round(j['roofline']['frac'],3), 'kms', round(j['roofline']['kernel_ms'],3))
"
done
