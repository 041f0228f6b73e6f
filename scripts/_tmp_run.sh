for mb in 12 24 48 100000; do
ALEXCAP_VOCAB_CHUNK_MB=$mb python bench.py --steps 30 --warmup 5 --no-cpu-baseline | python -c "import sys,json; j=json.loads(sys.stdin.read()); print($mb, j['ms_per_step'], j['value'])"
done
