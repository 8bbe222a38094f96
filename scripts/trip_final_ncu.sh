# refresh ncu evidence at HEAD: launch list of the default bench step + one --set full capture of the prefill attention kernel
mkdir -p gpurun_out
run() { name=$1; shift; timeout $1 "${@:2}" > gpurun_out/$name.log 2>&1; echo "== $name rc=$?"; tail -n ${TAILN:-6} gpurun_out/$name.log; }
run bench_plain 900 python bench.py --steps 2 --warmup 1 --sweep-prompts 0 --no-cpu-baseline
run ncu_launches 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_final.csv python bench.py --steps 2 --warmup 1 --sweep-prompts 0 --no-cpu-baseline
run prof_plain 300 python scripts/profile_attn.py 3
run ncu_attn 900 ncu --set full --clock-control none --import-source on -k regex:attn_prefill_sm100 -s 2 -c 1 -o gpurun_out/attn_v42 -f python scripts/profile_attn.py 2
ls -la gpurun_out | tail -12
