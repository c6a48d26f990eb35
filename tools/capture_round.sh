#!/usr/bin/env bash
# The capture set behind profiles/r<round><letter>_*: ONE call on one box.  Usage (on the GPU box, from the repo root):
#   bash tools/capture_round.sh <tag>        -> gpurun_out/<tag>/
# Every ncu pass runs only after the same command exited 0 without ncu; numbers printed under ncu are never bench values.
set -u
TAG=${1:-cap}
OUT=gpurun_out/$TAG
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1

run() {  # run <log> <command...>: log + exit code, never abort the set
    local log=$1; shift
    "$@" > "$OUT/$log" 2>&1
    echo "$log: exit $?" | tee -a "$OUT/status.txt"
}

nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > "$OUT/gpu.txt" 2>&1

# 1. the bench line of the contract and the reference arm
run bench_1gpu.log python bench.py
grep '^{' "$OUT/bench_1gpu.log" | tail -1 > "$OUT/bench_1gpu.json"
run bench_reference_arm.log python bench.py --impl reference --steps 3 --warmup 3
grep '^{' "$OUT/bench_reference_arm.log" | tail -1 > "$OUT/bench_reference_arm.json"

# 2. BASELINE.json configs on one GPU, both QP modes
run sweep_exact.log python tools/sweep_configs.py
run sweep_osqp.log python tools/sweep_configs.py --qp-mode osqp --dtypes fp64 --cases 1x32,64x32,512x32,4096x32,512x128
run sweep_osqp_limits.log python tools/sweep_configs.py --qp-mode osqp --limits --dtypes fp64 --cases 1x32,512x32
cat "$OUT"/sweep_exact.log "$OUT"/sweep_osqp.log "$OUT"/sweep_osqp_limits.log | grep '^{' > "$OUT/sweep_configs.jsonl"

# 3. cycle counters inside the kernels
: > "$OUT/phase_clocks.txt"
for b in 1 32 512 4096; do python tools/phase_profile.py --batch $b >> "$OUT/phase_clocks.txt" 2>&1; done
python tools/phase_profile.py --batch 512 --horizon 128 >> "$OUT/phase_clocks.txt" 2>&1
python tools/phase_profile.py --batch 512 --horizon 128 --dtype fp32 >> "$OUT/phase_clocks.txt" 2>&1
: > "$OUT/admm_phase_clocks.txt"
for b in 1 512 4096; do python tools/admm_profile.py --batch $b >> "$OUT/admm_phase_clocks.txt" 2>&1; done

# 4. launch list of a short bench run (after the same command without ncu)
SHORT="python bench.py --steps 3 --warmup 3 --latency-ticks 20 --no-cpu-baseline --config-ticks 0"
if $SHORT > "$OUT/bench_short.log" 2>&1; then
    ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file "$OUT/launches_bench.csv" $SHORT > "$OUT/ncu_launches.log" 2>&1
    echo "launches: exit $?" | tee -a "$OUT/status.txt"
fi

# 5. ncu --set full of the dominant kernels
full() {  # full <name> <kernel regex> <launch-skip> <command...>
    local name=$1 k=$2 skip=$3; shift 3
    if "$@" > "$OUT/$name.plain.log" 2>&1; then
        ncu --set full --clock-control none --import-source on -k "regex:$k" -s "$skip" -c 1 -f -o "$OUT/$name" "$@" > "$OUT/$name.ncu.log" 2>&1
        echo "$name: exit $?" | tee -a "$OUT/status.txt"
    else
        echo "$name: plain run failed" | tee -a "$OUT/status.txt"
    fi
}
full sqp_b4096 'sqp_kernel' 3 python tools/solve_loop.py --batch 4096 --ticks 6
full admm_b512 'admm_kernel' 3 python tools/admm_profile.py --batch 512
full sqp_b512_n128 'sqp_kernel' 3 python tools/solve_loop.py --batch 512 --horizon 128 --ticks 6
full cluster_b1 'sqp_cluster_kernel' 3 python tools/solve_loop.py --batch 1 --ticks 8
cat "$OUT/status.txt"
