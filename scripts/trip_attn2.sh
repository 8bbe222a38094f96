mkdir -p gpurun_out
run() { name=$1; shift; timeout $1 "${@:2}" > gpurun_out/$name.log 2>&1; echo "== $name rc=$?"; tail -n ${TAILN:-12} gpurun_out/$name.log; }
TAILN=3 run prof_attn 300 python scripts/profile_attn.py 10
for lib in cram_b200/_build/libcram_b200_*.so; do
  v=$(basename $lib .so | sed 's/libcram_b200_//')
  [ "$v" = "trace" ] && continue
  CRAM_B200_LIB=$lib TAILN=3 run prof_attn_$v 300 python scripts/profile_attn.py 10
done
