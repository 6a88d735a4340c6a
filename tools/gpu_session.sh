#!/usr/bin/env bash
# One GPU-box visit, many measurements: every gpurun call pays ~20 s of fixed overhead, so batch.
#   gpurun --timeout 1500 -- 'bash tools/gpu_session.sh tests bench ops logistic tour'
# Stages write under gpurun_out/session/ (small text / JSON only; ncu reports are exported to CSV
# on the box and deleted, the 64 MiB copy-back limit is easy to hit with .ncu-rep files).
set -u
out=gpurun_out/session
mkdir -p "$out"
for stage in "$@"; do
  echo "=== $stage ==="
  case "$stage" in
    tests)    timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee "$out/tests.txt" ;;
    bench)    timeout 300 python bench.py --steps 20 --warmup 3 2>"$out/bench.err" | tail -1 | tee "$out/bench_1gpu.json" | cut -c1-300 ;;
    smoke)    python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE_OK')" 2>&1 | tail -1 ;;
    ops)      timeout 300 python tools/perf_ops.py all > "$out/perf_ops_b200.json" 2>"$out/perf_ops.err"; tail -c 400 "$out/perf_ops_b200.json" ;;
    logistic) timeout 300 python benchmarks/logistic_scaled_bench.py --workers 512 --dim 8192 --samples 1024 --iters 15 --rho 3 --quiet 2>&1 | tail -1 | tee "$out/logistic_512x8192.json" | cut -c1-300 ;;
    gram)     timeout 120 python tools/gram_kseg_probe.py 2>&1 | tail -6 | tee "$out/gram_probe.txt" ;;
    inverse)  timeout 120 python tools/inverse_breakdown.py >/dev/null 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file "$out/inverse_launches.csv" python tools/inverse_breakdown.py >/dev/null 2>&1; tail -2 "$out/inverse_launches.csv" ;;
    tour)     timeout 200 python tools/ncu_tour.py 2>&1 | tail -1 && timeout 500 ncu --section SpeedOfLight --section LaunchStats --section Occupancy --section MemoryWorkloadAnalysis --clock-control none -k regex:'^(logistic_|symv_grad|shared_|own_exchange|dgd_update|dualavg|star_|find_path|rechain|gadmm_persistent|gadmm_half|gadmm_finish|gadmm_dual)' -c 160 -f -o /tmp/kernel_tour python tools/ncu_tour.py >/dev/null 2>&1; ncu -i /tmp/kernel_tour.ncu-rep --page raw --csv > "$out/kernel_tour_raw.csv" 2>/dev/null; rm -f /tmp/kernel_tour.ncu-rep; wc -c "$out/kernel_tour_raw.csv" ;;
    mp)       n=$(nvidia-smi -L | wc -l); timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port 29581 tests/mp_gadmm_check.py > "$out/mp_${n}gpu.log" 2>&1; grep -n "OK\|Error\|assert" "$out/mp_${n}gpu.log" | head -8 ;;
    mpbench)  n=$(nvidia-smi -L | wc -l); timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port 29582 bench.py --gpus "$n" --steps 20 --warmup 3 2>"$out/bench_${n}gpu.err" | tail -1 | tee "$out/bench_${n}gpu.json" | cut -c1-300 ;;
    *)        echo "unknown stage $stage" ;;
  esac
done
du -sh gpurun_out | tail -1
