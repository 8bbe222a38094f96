mkdir -p gpurun_out
python scripts/profile_decode.py 5 0 > gpurun_out/prof_decode_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:attn_decode_mma -s 30 -c 2 -o gpurun_out/decode_full -f \
  python scripts/profile_decode.py 3 0 > gpurun_out/ncu_decode.log 2>&1; echo "ncu decode rc=$?"
python bench.py --workload decode --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/bench_decode_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --graph-profiling node -s 4000 -c 700 --csv --log-file gpurun_out/launches_decode.csv \
  python bench.py --workload decode --steps 1 --warmup 1 --no-cpu-baseline --new-tokens 8 > gpurun_out/ncu_launches_decode.log 2>&1; echo "ncu launches decode rc=$?"
tail -3 gpurun_out/prof_decode_plain.log; ls -la gpurun_out | tail -8
