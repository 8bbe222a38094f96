# full GPU trip: every -m gpu test, smoke, default bench, reference arm, launch list, ncu --set full of attention
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
run() { name=$1; shift; timeout $1 "${@:2}" > gpurun_out/$name.log 2>&1; echo "== $name rc=$?"; tail -n ${TAILN:-12} gpurun_out/$name.log; }
TAILN=25 run t_gpu 1500 python -m pytest tests -m gpu -q -x
run smoke 600 python __graft_entry__.py --smoke
TAILN=5 run bench 1500 python bench.py
TAILN=5 run bench_ref 900 python bench.py --impl reference --steps 2 --warmup 1
run prof_attn 300 python scripts/profile_attn.py 10
python bench.py --steps 1 --warmup 1 --sweep-prompts 0 --no-cpu-baseline > gpurun_out/bench_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 250 -c 800 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 1 --warmup 1 --sweep-prompts 0 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:attn_prefill_sm100 -s 2 -c 1 -o gpurun_out/attn_full -f \
  python scripts/profile_attn.py 2 > gpurun_out/ncu_attn.log 2>&1; echo "ncu attn rc=$?"
ls -la gpurun_out | head -30
