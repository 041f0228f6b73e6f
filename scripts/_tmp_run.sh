python -m pytest tests -m gpu -x -q 2>&1 | grep -E "passed|failed|Error|bad|assert" | tail -6
python bench.py --steps 30 --warmup 5 --no-cpu-baseline | python -c "import sys,json; j=json.loads(sys.stdin.read()); print(j['ms_per_step'], j['value'], j['e2e']['value'], j['loss'], j['roofline']['ms_per_launch'], j['roofline']['backward']['ms_per_launch'])"
