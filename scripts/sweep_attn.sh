# ring geometry sweep of the streaming attention kernels (graph-timed): stage rows x ring depth
python scripts/bench_attn.py 2>&1 | tail -1
for rows in 8 16; do for nst in 2 4 6 8 12; do
echo "rows=$rows nst=$nst: $(ALEXCAP_ATT_ROWS=$rows ALEXCAP_ATT_NST=$nst timeout 60 python scripts/bench_attn.py 2>&1 | tail -1)"
done; done
