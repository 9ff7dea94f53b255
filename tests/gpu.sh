#!/bin/bash
# One parameterised driver for everything that runs on the GPU box (replaces the per-session gpu_run*.sh scripts):
#   gpurun --timeout 1800 -- 'bash tests/gpu.sh TAG step[:arg] step[:arg] ...'
# Every step has its own timeout, writes gpurun_out/TAG_<step>.out and appends its exit code + tail to gpurun_out/TAG.log;
# a failing step does not stop the later ones. Steps:
#   box                      GPU / host facts
#   smoke                    __graft_entry__.smoke()
#   pytest[:files / -k expr] python -m pytest -m gpu -q <files ... | tests>   (give the directory yourself with -k: "tests -k expr")
#   pytestx[:...]            the same with -x
#   pytestmon                the whole suite with -x --durations=30 and a memory monitor (TAG_pytestmon<n>.mem / .dmesg)
#   sweep:"runs"             tests/sweep_variants.py --scale 1.0 --runs <runs> (variant libraries from tests/build_variants.sh)
#   sweepw:"workload scale runs"
#   bench[:args]             bench.py --steps 5 --warmup 3 <args>   (default args: --no-e2e --no-cpu)
#   benchfull                bench.py with its defaults (the driver's line)
#   benchref                 bench.py --impl reference --steps 3 --warmup 1
#   launches[:bench args]    ncu launch list (gpu__time_duration) of bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-verify
#   ncufull:"regex skip count [workload scale]"   ncu --set full of tests/profile_join.py, report -> gpurun_out/TAG_ncu_<n>.ncu-rep
#   prof:"args"              tests/profile_join.py --joins 3 <args> (per-phase timings, e.g. "--scale 1.0 --sharded")
#   micro:name               tests/micro/<name> (prebuilt binary)
#   sharded[:args]           torchrun -n $NGPU tests/sharded_check.py --backend nccl <args>
#   benchn[:args]            torchrun -n $NGPU bench.py --gpus $NGPU --steps 5 --warmup 3 <args>
#   py:"script args"         python <script args>
cd "$(dirname "$0")/.."
TAG=$1; shift
mkdir -p gpurun_out
L=gpurun_out/$TAG.log; : > $L
NGPU=${NGPU:-$(nvidia-smi -L | wc -l)}
n=0
run() {  # run <name> <timeout s> <command...>
  local name=$1 t=$2; shift 2
  local out="gpurun_out/${TAG}_${name}.out"
  ( time timeout "$t" "$@" > "$out" 2>&1 ) 2>> $L
  local rc=$?
  echo "== $name rc=$rc" | tee -a $L
  tail -${TAIL:-6} "$out" | cut -c1-${CUT:-600} | tee -a $L
}
TORCHRUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NGPU --master-addr 127.0.0.1 --master-port"
for spec in "$@"; do
  step=${spec%%:*}; arg=""; [ "$spec" != "$step" ] && arg=${spec#*:}
  n=$((n + 1))
  case $step in
    box) run box 30 bash -c 'nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv; nvidia-smi topo -m 2>/dev/null | head -12; nproc; free -g | head -2' ;;
    smoke) run smoke 300 python __graft_entry__.py smoke ;;
    pytest) run pytest$n 1500 python -m pytest -m gpu -q ${arg:-tests} ;;
    pytestx) run pytest$n 1500 python -m pytest -m gpu -q -x ${arg:-tests} ;;
    pytestmon) # the whole suite as the driver runs it, with the process's memory sampled every 5 s and the slowest tests listed
             ( while sleep 5; do p=$(pgrep -n -f "pytest tests -m gpu" ); [ -n "$p" ] && echo "$(date +%s) rss_kb=$(awk '/VmRSS/{print $2}' /proc/$p/status 2>/dev/null) $(free -m | awk '/Mem:/{print "used_mb="$3" free_mb="$4}') $(nvidia-smi --query-gpu=memory.used --format=csv,noheader)"; done > gpurun_out/${TAG}_pytestmon$n.mem 2>&1 ) & mon=$!
             run pytestmon$n 1500 python -m pytest tests -m gpu -q -x --durations=30 $arg
             kill $mon 2>/dev/null; (dmesg 2>/dev/null | tail -20) > gpurun_out/${TAG}_pytestmon$n.dmesg 2>&1 ;;
    sweep) run sweep$n 600 python tests/sweep_variants.py --scale 1.0 --runs $arg --out gpurun_out/${TAG}_sweep$n.json ;;
    sweepw) set -- $arg; w=$1; sc=$2; shift 2; run sweep$n 600 python tests/sweep_variants.py --workload $w --scale $sc --joins 3 --runs "$@" --out gpurun_out/${TAG}_sweep$n.json ;;
    bench) run bench$n 600 python bench.py --steps 5 --warmup 3 ${arg:---no-e2e --no-cpu} ;;
    benchfull) run benchfull 900 python bench.py ;;
    benchref) run benchref 600 python bench.py --impl reference --steps 3 --warmup 1 ;;
    launches) run launches$n 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_ -c 400 --csv --log-file gpurun_out/${TAG}_launches$n.csv \
                python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-verify $arg ;;
    ncufull) set -- $arg; rx=$1; skip=${2:-0}; cnt=${3:-4}; w=${4:-c2}; sc=${5:-1.0}
             run ncu$n 1200 ncu --set full --clock-control none --import-source on -k "regex:$rx" -s $skip -c $cnt -o gpurun_out/${TAG}_ncu$n -f \
                python tests/profile_join.py --workload $w --scale $sc --joins 2 $6
             # what comes back is text (gpurun_out/ may hold 64 MiB): the raw metrics and the per-line source counters
             ncu -i gpurun_out/${TAG}_ncu$n.ncu-rep --page raw --csv > gpurun_out/${TAG}_ncu$n.raw.csv 2>/dev/null
             ncu -i gpurun_out/${TAG}_ncu$n.ncu-rep --page source --csv --print-source sass,cuda 2>/dev/null | gzip -9 > gpurun_out/${TAG}_ncu$n.source.csv.gz
             ncu -i gpurun_out/${TAG}_ncu$n.ncu-rep --page details --csv 2>/dev/null | gzip -9 > gpurun_out/${TAG}_ncu$n.details.csv.gz
             [ $(stat -c %s gpurun_out/${TAG}_ncu$n.ncu-rep) -gt 12000000 ] && rm -f gpurun_out/${TAG}_ncu$n.ncu-rep ;;
    prof) run prof$n 300 python tests/profile_join.py --joins 3 $arg ;;
    micro) run micro_$arg 120 tests/micro/$arg ;;
    sharded) run sharded$n 900 $TORCHRUN $((29500 + RANDOM % 400)) tests/sharded_check.py --backend nccl $arg ;;
    benchn) run benchn$n 900 $TORCHRUN $((29500 + RANDOM % 400)) bench.py --gpus $NGPU --steps 5 --warmup 3 ${arg:---no-e2e --no-cpu} ;;
    py) run py$n 900 python $arg ;;
    *) echo "unknown step $step" | tee -a $L ;;
  esac
done
