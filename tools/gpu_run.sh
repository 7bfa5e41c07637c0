#!/bin/bash
# The one runner.  On the GPU box (called through gpurun from the build container):
#   bash tools/gpu_run.sh <tag> <mode> [args...]          results under gpurun_out/<tag>/
# In the build container (no GPU):
#   bash tools/gpu_run.sh <tag> build-rev <git-rev>        full library build of a git revision for A/B runs ->
#                                                          radiative-transfer_b200/csrc/build/ab/<tag>.so (select it with RT_CUDA_LIB)
# modes (several may be chained with "--"):
#   tests [pytest args]     pytest -m gpu (default: whole suite, -x)
#   bench [bench.py args]   one bench.py line -> bench_<n>.json
#   ab <name=ENV=VAL,...>.. device-timed A/B of library/env variants on one workload (WL=c4|wide|..., default c4)
#   launches [bench args]   ncu launch list (gpu__time_duration.sum) of a short bench run -> launches.csv
#   ncu <kernel-regex> [bench args]   one --set full capture of the first matching launch -> full.ncu-rep + digest
TAG=$1; shift
if [ "$1" == "build-rev" ]; then
  set -e
  AB=radiative-transfer_b200/csrc/build/ab; D=$AB/${TAG}_tree
  rm -rf $D; mkdir -p $D
  git archive $2 radiative-transfer_b200/csrc include | tar -x -C $D
  make -C $D/radiative-transfer_b200/csrc > /dev/null
  cp $D/radiative-transfer_b200/librt_cuda.so $AB/$TAG.so
  rm -rf $D
  echo built $AB/$TAG.so
  exit 0
fi
OUT=gpurun_out/$TAG; mkdir -p $OUT
n=0
run_mode() {
  local mode=$1; shift
  case $mode in
  tests)
    local args=("$@"); [ ${#args[@]} -eq 0 ] && args=(tests -x)
    timeout ${TEST_TIMEOUT:-1500} python -m pytest "${args[@]}" -m gpu -q --durations=8 > $OUT/pytest_gpu_$n.log 2>&1
    echo "pytest exit $?"; tail -${TAIL:-25} $OUT/pytest_gpu_$n.log ;;
  bench)
    timeout ${BENCH_TIMEOUT:-600} python bench.py "$@" > $OUT/bench_$n.json 2> $OUT/bench_$n.err; echo "bench exit $?"
    tail -3 $OUT/bench_$n.err
    python - $OUT/bench_$n.json <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1])); r = d["roofline"]
    print(f"value {d['value']:.4e} ms/step {d['ms_per_step']:.2f} frac {r['frac']:.3f} launch_ms {r['avg_launch_ms']:.2f} e2e {d['e2e']['value'] if d.get('e2e') else None} batches {d['config'].get('batches')} clocks {d['clocks']}")
    for k in ("c5", "temperature_coupled"):
        if d.get(k): print(k, json.dumps(d[k])[:600])
except Exception as e:
    print("bench line unreadable:", e)
PY
    ;;
  ab)
    for spec in "$@"; do
      name=${spec%%=*}; envs=${spec#*=}; [ "$envs" == "$spec" ] && envs=""
      ( IFS=','; for kv in $envs; do export "$kv"; done
        timeout 240 python bench.py --workload ${WL:-c4} --steps ${STEPS:-3} --warmup ${WARMUP:-3} --no-e2e --no-cpu-baseline ${BENCH_ARGS} > $OUT/ab_$name.json 2> $OUT/ab_$name.err || tail -3 $OUT/ab_$name.err )
      python - $OUT/ab_$name.json $name <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1])); r = d["roofline"]
    print(f"{sys.argv[2]:14s} value {d['value']:.4e} ms/step {d['ms_per_step']:.2f} frac {r['frac']:.3f} launch_ms {r['avg_launch_ms']:.3f} grid {d['sweep_grid']} batches {d['config'].get('batches')} MHz {d['clocks'].get('sm_mhz')} W {d['clocks'].get('power_w_max')} {d['clocks'].get('reasons')}")
except Exception as e:
    print(sys.argv[2], "failed", e)
PY
    done ;;
  launches)
    timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c ${NCU_COUNT:-400} --csv --log-file $OUT/launches.csv \
        python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-extras "$@" > $OUT/launches_run.log 2>&1
    echo "ncu launch list exit $?"; grep -c . $OUT/launches.csv ;;
  ncu)
    local re=$1; shift
    timeout 1500 ncu --set full --import-source on --clock-control none -k regex:$re -s ${NCU_SKIP:-2} -c 1 -f -o $OUT/full_$n \
        python bench.py --steps 1 --warmup 2 --no-e2e --no-cpu-baseline --no-extras "$@" > $OUT/ncu_run_$n.log 2>&1
    echo "ncu full exit $?"; ls -la $OUT/full_$n.ncu-rep 2>/dev/null
    ncu -i $OUT/full_$n.ncu-rep --page raw --csv > $OUT/full_${n}_raw.csv 2>/dev/null ;;
  multigpu)
    # everything that needs N GPUs of one box, N = $1: host-drain ceiling (bound / unbound), bench under torchrun, the
    # single-process path of the C ABI
    local N=$1; shift
    local TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
    for k in 1 2 4 8; do
      [ $k -gt $N ] && continue
      for b in 1 0; do
        RT_NUMA_BIND=$b timeout 120 $TR --nproc-per-node $k --master-port $((29600 + k)) tools/d2h_ceiling.py 2>/dev/null | grep '^{' > $OUT/d2h_n${k}_bind$b.json
        python -c "import json; d=json.load(open('$OUT/d2h_n${k}_bind$b.json')); print('d2h n=$k bind=$b per-gpu min/max %.1f/%.1f total %.1f GB/s' % (d['GBps_per_gpu_min'], d['GBps_per_gpu_max'], d['GBps_total']), d['numa_bind_rank0'])" 2>&1 | tail -1
      done
    done
    timeout 900 $TR --nproc-per-node $N --master-port 29700 bench.py --gpus $N --steps 5 --warmup 3 > $OUT/bench_n$N.json 2> $OUT/bench_n$N.err; echo "bench n=$N exit $?"
    RT_NUMA_BIND=0 timeout 600 $TR --nproc-per-node $N --master-port 29701 bench.py --gpus $N --steps 5 --warmup 3 --no-extras > $OUT/bench_n${N}_nobind.json 2> $OUT/bench_n${N}_nobind.err; echo "bench n=$N unbound exit $?"
    timeout 600 python bench.py --in-process-gpus $N --steps 5 --warmup 3 > $OUT/bench_inprocess_n$N.json 2> $OUT/bench_inprocess_n$N.err; echo "in-process n=$N exit $?"
    timeout 600 python bench.py --in-process-gpus $N --workload c5r --cv 1 --steps 10 --warmup 3 > $OUT/bench_inprocess_cv_n$N.json 2> $OUT/bench_inprocess_cv_n$N.err; echo "in-process cv n=$N exit $?"
    python - $OUT $N <<'PY'
import json, sys
out, N = sys.argv[1], sys.argv[2]
for f in (f"bench_n{N}.json", f"bench_n{N}_nobind.json", f"bench_inprocess_n{N}.json", f"bench_inprocess_cv_n{N}.json"):
    try:
        d = json.load(open(f"{out}/{f}"))
        e = d.get("e2e") or {}
        print(f, f"value {d['value']:.4e} per_gpu {d['per_gpu_value']:.4e} ms/step {d['ms_per_step']:.2f} e2e {e.get('value')} e2e/value {(e.get('value') or 0) / d['value']:.3f}")
        tc = d.get("temperature_coupled")
        if tc: print("   tc", {k: tc[k] for k in ("value", "ms_per_step", "exchange_ms_per_step", "allreduce", "parity_ok", "T_identical_on_all_ranks")})
    except Exception as ex:
        print(f, "unreadable:", ex)
PY
    ;;
  *) echo "unknown mode $mode"; return 2 ;;
  esac
}
args=()
for a in "$@"; do
  if [ "$a" == "--" ]; then run_mode "${args[@]}"; n=$((n+1)); args=(); else args+=("$a"); fi
done
[ ${#args[@]} -gt 0 ] && run_mode "${args[@]}"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv,noheader > $OUT/gpu.txt 2>&1
exit 0
