mkdir -p gpurun_out
python scripts/profile_attn.py 3 > gpurun_out/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:attn_prefill_sm100 -s 2 -c 1 -o gpurun_out/attn_v2 -f \
  python scripts/profile_attn.py 2 > gpurun_out/ncu_attn.log 2>&1; echo "ncu attn rc=$?"
tail -3 gpurun_out/prof_plain.log
