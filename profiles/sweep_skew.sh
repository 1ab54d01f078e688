for cfg in "0 0" "1 1500" "1 2500" "1 4000" "2 0"; do set -- $cfg
ADSBCU_SKEW=$1 ADSBCU_SKEW_NS=$2 python bench.py --no-cpu-baseline --steps 10 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('skew $1 ns $2', round(d['value']), round(d['e2e']['value']), round(d['roofline']['frac'],4), d['breakdown_ms'])"
done
