#!/bin/bash
# One gpurun call: GPU parity tests, perf probe, bench (ours + reference arm), ncu launch list + full capture.
# Usage (from the repo root on the GPU box): bash tools/gpu_check.sh [tests|probe|star|chain|stress|bench|configs|ncu|kernels|lazy|schedule|e2e|sanitize ...]
set -u
mkdir -p gpurun_out
WHAT="${*:-tests probe bench ncu kernels lazy}"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
lscpu | head -20 > gpurun_out/lscpu.txt 2>&1
for w in $WHAT; do
case $w in
tests)
  timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
  tail -5 gpurun_out/pytest_gpu.log ;;
probe)
  timeout 600 python tools/perf_probe.py > gpurun_out/probe.log 2>&1; echo "probe rc=$?"; cat gpurun_out/probe.log ;;
bench)
  timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
  timeout 600 python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json ;;
ncu)
  # launch list of the bench command, then one full capture each of a hub-bound and a hub-free launch of the span kernel
  CMD="python bench.py --edges 400000 --steps 1 --warmup 3 --no-cpu --e2e-batches 10 --no-stable --no-numba"
  timeout 600 $CMD > gpurun_out/ncu_plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
  echo "ncu launches rc=$?"
  for m in hub flat; do
    timeout 300 python tools/ncu_targets.py $m > gpurun_out/ncu_target_$m.log 2>&1 &&
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:tppr_span_kernel -s 3 -c 2 -f -o gpurun_out/prof_$m python tools/ncu_targets.py $m > gpurun_out/ncu_full_$m.log 2>&1
    echo "ncu full $m rc=$?"; tail -2 gpurun_out/ncu_full_$m.log
  done ;;
kernels)
  timeout 600 python tools/perf_kernels.py > gpurun_out/kernels.log 2>&1; echo "kernels rc=$?"; cat gpurun_out/kernels.log
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gather_fwd8|prune_topk" -c 6 -f -o gpurun_out/prof_kernels python tools/perf_kernels.py > gpurun_out/ncu_kernels.log 2>&1
  echo "ncu kernels rc=$?" ;;
lazy)
  # lazy memory update (level 2): microbench + launch list of the selection / lazy gather kernels
  timeout 300 python tools/perf_lazy.py > gpurun_out/perf_lazy.log 2>&1; echo "perf_lazy rc=$?"; cat gpurun_out/perf_lazy.log
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"pending_|gather_" -c 60 --csv --log-file gpurun_out/launches_lazy.csv python tools/perf_lazy.py > gpurun_out/ncu_lazy.log 2>&1
  echo "ncu lazy rc=$?" ;;
sanitize)
  # compute-sanitizer over smoke() (streaming path against the oracle): memcheck, racecheck (shared memory), synccheck
  for tool in memcheck racecheck synccheck; do
    timeout 900 compute-sanitizer --tool $tool --log-file gpurun_out/sanitizer_$tool.log python __graft_entry__.py smoke > gpurun_out/sanitizer_$tool.out 2>&1
    echo "sanitizer $tool rc=$?"; tail -3 gpurun_out/sanitizer_$tool.log; tail -1 gpurun_out/sanitizer_$tool.out
  done ;;
star)
  ZEBRA_TPPR_LIB=zebra_b200/lib/libzebra_tppr_prof.so timeout 600 python tools/perf_probe.py --star > gpurun_out/star_prof.log 2>&1; echo "star prof rc=$?"; cat gpurun_out/star_prof.log
  timeout 600 python tools/perf_probe.py --star > gpurun_out/star.log 2>&1; echo "star rc=$?"; cat gpurun_out/star.log ;;
chain)
  # per-phase cycles of the hottest chain (profile build) and plain timings on a Wiki-Talk-shaped prefix
  ZEBRA_TPPR_LIB=zebra_b200/lib/libzebra_tppr_prof.so timeout 600 python tools/perf_chain.py > gpurun_out/chain_prof.log 2>&1; echo "chain prof rc=$?"; cat gpurun_out/chain_prof.log
  timeout 600 python tools/perf_chain.py > gpurun_out/chain.log 2>&1; echo "chain rc=$?"; cat gpurun_out/chain.log ;;
stress)
  timeout 900 python tools/stress_chains.py 60 > gpurun_out/stress_chains.log 2>&1; echo "stress rc=$?"; tail -3 gpurun_out/stress_chains.log ;;
configs)
  for w in wikipedia reddit mooc superuser; do
    timeout 600 python bench.py --workload $w > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"; done ;;
schedule)
  # announced passes: per-batch cost per call against an announced pass, four stream shapes; the
  # allocation-free pruning entry point in a process of its own
  for w in reddit:120000 wikitalk:400000 wikipedia:157474 superuser:400000; do
    timeout 300 python tools/perf_schedule.py --workload ${w%%:*} --edges ${w##*:}
  done > gpurun_out/perf_schedule.log 2>&1; cat gpurun_out/perf_schedule.log
  timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "no_alloc alone rc=$?" ;;
e2e)
  # the reference's train.py, reference arm against levels 1 and 2 (three configs would take ~6 GPU-minutes: one here)
  timeout 1500 python tools/e2e_train.py --ref baseline/_ref --name ${E2E_NAME:-wikipedia} --epochs 2 --levels ref 1 2 3 --out gpurun_out/e2e 2>&1 | tail -5 ;;
esac
done
