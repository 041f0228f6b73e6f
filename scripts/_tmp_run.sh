for f in 0 0.45 0.6 0.75; do
echo "frac=$f: $(ALEXCAP_ATT_L2_FRAC=$f timeout 200 python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null < /dev/null | cut -c50-175)"
done
