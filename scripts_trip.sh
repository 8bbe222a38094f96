mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
run() { name=$1; shift; timeout $1 "${@:2}" > gpurun_out/$name.log 2>&1; echo "== $name rc=$?"; tail -n 12 gpurun_out/$name.log; }
run t_span 400 python -m pytest tests/test_span_index_gpu.py -q -s
run t_glue 400 python -m pytest tests/test_glue_gpu.py -q -s
run t_attn_generic 600 python -m pytest tests/test_attention_gpu.py -q -s -k generic_vs
run t_attn_sm100 900 python -m pytest tests/test_attention_gpu.py -q -s -k "not generic_vs"
run t_decode 600 python -m pytest tests/test_decode_gpu.py -q -s
run t_engine 900 python -m pytest tests/test_engine_gpu.py -q -s
run t_sweep 900 python -m pytest tests/test_sweep_gpu.py -q -s
run bench1 1200 python bench.py --steps 3 --warmup 1 --sweep-prompts 0 --no-cpu-baseline
