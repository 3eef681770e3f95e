for n in 1 8 8 16 4; do
  python bench.py --no-cpu-baseline --steps 200 --event-every $n 2>/dev/null > /tmp/o.json
  python -c "import json; d=json.load(open('/tmp/o.json')); print($n, round(d['value']), round(d['ms_per_step']*1000,2), d['kernel_ms'], round(d['roofline']['frac'],4))"
done
