mkdir -p gpurun_out
python scripts/profile_linear.py 10 > gpurun_out/prof_linear.log 2>&1; tail -10 gpurun_out/prof_linear.log
python scripts/profile_linear.py 10 chain > gpurun_out/prof_chain.log 2>&1; tail -8 gpurun_out/prof_chain.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:linear_decode -c 8 -o gpurun_out/linear_r1 -f python scripts/profile_linear.py 1 eager > gpurun_out/ncu_linear.log 2>&1; echo "ncu rc=$?"
python bench.py --workload decode --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/bench_decode_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --graph-profiling node -s 4000 -c 700 --csv --log-file gpurun_out/launches_decode.csv \
  python bench.py --workload decode --steps 1 --warmup 1 --no-cpu-baseline --new-tokens 8 > gpurun_out/ncu_launches_decode.log 2>&1; echo "ncu launches decode rc=$?"
