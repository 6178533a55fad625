for cfg in "${@:-16 6}"; do set -- $cfg; echo "F=$1 D=$2"; python bench.py --steps 1920 --warmup 64 --frames-per-batch $1 --in-flight $2 --no-sweep --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('  value',d['value'],'e2e',d['e2e']['value'],'kernel_ms',d['roofline']['kernel_ms'])
    elif 'rror' in l: print(l.strip())
"; done
