mkdir -p gpurun_out
run() { name=$1; shift; timeout $1 "${@:2}" > gpurun_out/$name.log 2>&1; echo "== $name rc=$?"; tail -n ${TAILN:-12} gpurun_out/$name.log; }
run t_span 400 python -m pytest tests/test_span_index_gpu.py -q -s
TAILN=40 run t_attn_generic 600 python -m pytest tests/test_attention_gpu.py -q -s -k generic_vs
TAILN=60 run t_attn_sm100 900 python -m pytest tests/test_attention_gpu.py -q -s -k "not generic_vs"
TAILN=30 run t_decode 600 python -m pytest tests/test_decode_gpu.py -q -s
run t_engine 900 python -m pytest tests/test_engine_gpu.py -q -s
run prof_attn 300 python scripts/profile_attn.py 5
python bench.py --steps 1 --warmup 1 --sweep-prompts 0 --no-cpu-baseline > gpurun_out/bench_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 250 -c 800 --csv --log-file gpurun_out/launches_r1.csv \
  python bench.py --steps 1 --warmup 1 --sweep-prompts 0 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
python scripts/profile_attn.py 2 > gpurun_out/prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:attn_prefill_sm100 -s 2 -c 2 -o gpurun_out/attn_r1 -f \
  python scripts/profile_attn.py 2 > gpurun_out/ncu_attn.log 2>&1; echo "ncu attn rc=$?"
ls -la gpurun_out | head -30
