for t in 2 4 8; do
ADSBCU_HOST_TPS=$t python bench.py --no-cpu-baseline --steps 10 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('host_tps $t', round(d['value']), round(d['e2e']['value']), round(d['roofline']['frac'],4), d['breakdown_ms'], round(d['pcie']['h2d_ceiling_msamples_s']))"
done
