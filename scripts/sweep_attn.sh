python -m pytest tests -m gpu -x -q 2>&1 | grep -E "passed|failed|Error|bad" | tail -8
python scripts/bench_attn.py
for rows in 8 16; do for nst in 3 5; do
ALEXCAP_ATT_ROWS=$rows ALEXCAP_ATT_NST=$nst timeout 60 python scripts/bench_attn.py 2>&1 | tail -1
done; done
