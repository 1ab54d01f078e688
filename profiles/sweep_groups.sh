# timing experiment (results incomplete for skew >= 10): how well do the 4 groups of an SM overlap?
for k in 11 12 13 14; do
ADSBCU_SKEW=$k python bench.py --no-cpu-baseline --steps 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('active groups', $k-10, d['breakdown_ms'])"
done
