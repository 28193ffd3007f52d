#!/bin/bash
# One parametrised GPU checkpoint script (run under gpurun):
#   gpurun --timeout 1500 -- 'bash scripts/gpu_checkpoint.sh <tag> <step> [<step> ...]'
# steps (each writes gpurun_out/<tag>_<step>.*; a failing step does not stop the later ones):
#   tests            python -m pytest tests -m gpu -x -q
#   tests:<expr>     the same with -k <expr>
#   testsall         the whole GPU suite without -x (every failure listed)
#   smoke            __graft_entry__.smoke()
#   bench            python bench.py --steps 3 --warmup 3            (default line with every extra)
#   benchq           python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline
#   ref              python bench.py --impl reference --steps 2 --warmup 1
#   train            python bench.py --workload train --steps 20 --warmup 5
#   latency|sweep|libero|eager    python bench.py --workload <step>
#   launches         ncu launch list of 2 denoise steps at 65536 rows (one lane): per-kernel device times
#   launches-train   ncu launch list of one serialised training step
#   ncu:<regex>      ncu --set full of 3 launches of the kernels matching <regex> in the same 2-step command
#   sanitize:<tool>  compute-sanitizer --tool <memcheck|racecheck|synccheck|initcheck> on the reduced test subset
# Per B200_PROFILING.md use at most ONE ncu / compute-sanitizer step per call.
set -u
tag=$1; shift
mkdir -p gpurun_out
O=gpurun_out/${tag}
SHORT="python bench.py --rows 65536 --denoise-steps 2 --steps 1 --warmup 1 --no-cpu-baseline --no-extras"
# reduced subset for the sanitizers: one pair-GEMM case per epilogue kind, the chain kernel, quad clusters, the
# LayerNorm pre-phase barrier, the fused final layer, 333-row training, the decoder tail
SUBSET='test_umma_pair_gemm_layouts and (300 or 129) or test_pair_gemm_bf16_output_epilogues and 256 or test_ragged_rows_vs_oracle or test_latency_path_equals_chunked_path and (16-1.0 or 27-1.0) or test_chain_kernel_cluster_staging or test_lnmod_prephase and 300 or test_fused_final_layer_ddpm or test_ragged_masked_rows_vs_oracle or test_decoder_tail_rows'
for step in "$@"; do
  name=${step%%:*}; arg=${step#*:}; [ "$arg" = "$step" ] && arg=""
  echo "=== [$tag] $step ($(date +%T))"
  case $name in
    tests)
      if [ -n "$arg" ]; then timeout 1500 python -m pytest tests -m gpu -x -q -k "$arg" > ${O}_tests.log 2>&1
      else timeout 1500 python -m pytest tests -m gpu -x -q > ${O}_tests.log 2>&1; fi
      echo "rc=$?"; tail -n 15 ${O}_tests.log ;;
    testsall) timeout 2400 python -m pytest tests -m gpu -q > ${O}_tests.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed" ${O}_tests.log | tail -n 30 ;;
    smoke) timeout 600 python __graft_entry__.py smoke > ${O}_smoke.log 2>&1; echo "rc=$?"; tail -n 4 ${O}_smoke.log ;;
    bench) timeout 900 python bench.py --steps 3 --warmup 3 > ${O}_bench.json 2> ${O}_bench.err; echo "rc=$?"; cut -c1-1500 ${O}_bench.json; tail -n 3 ${O}_bench.err ;;
    benchq) timeout 600 python bench.py --steps 3 --warmup 3 --no-extras --no-cpu-baseline > ${O}_benchq.json 2> ${O}_benchq.err; echo "rc=$?"; cut -c1-1200 ${O}_benchq.json; tail -n 3 ${O}_benchq.err ;;
    ref) timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > ${O}_ref.json 2> ${O}_ref.err; echo "rc=$?"; cut -c1-600 ${O}_ref.json ;;
    train) timeout 600 python bench.py --workload train --steps 20 --warmup 5 > ${O}_train.json 2> ${O}_train.err; echo "rc=$?"; cut -c1-900 ${O}_train.json; tail -n 3 ${O}_train.err ;;
    latency|sweep|libero|eager) timeout 900 python bench.py --workload $name > ${O}_$name.json 2> ${O}_$name.err; echo "rc=$?"; cut -c1-1500 ${O}_$name.json; tail -n 3 ${O}_$name.err ;;
    launches)
      UVA_OVERLAP=0 $SHORT > ${O}_launches_plain.log 2>&1 &&
      UVA_OVERLAP=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file ${O}_launches.csv $SHORT > ${O}_launches_ncu.log 2>&1
      echo "rc=$?"; python scripts/agg_launches.py ${O}_launches.csv > ${O}_launches_agg.txt 2>&1; head -n 30 ${O}_launches_agg.txt ;;
    launches-train)
      CMD="python bench.py --workload train --steps 1 --warmup 2"
      UVA_ACT_OVERLAP=0 UVA_TRAIN_GRAPH=0 $CMD > ${O}_ltrain_plain.log 2>&1 &&
      UVA_ACT_OVERLAP=0 UVA_TRAIN_GRAPH=0 timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file ${O}_launches_train.csv $CMD > ${O}_ltrain_ncu.log 2>&1
      echo "rc=$?"; python scripts/agg_launches.py ${O}_launches_train.csv > ${O}_launches_train_agg.txt 2>&1; head -n 40 ${O}_launches_train_agg.txt ;;
    ncu)
      UVA_OVERLAP=0 $SHORT > ${O}_ncu_plain.log 2>&1 &&
      UVA_OVERLAP=0 timeout 1200 ncu --set full --clock-control none --import-source on -k "regex:$arg" -s ${NCU_SKIP:-6} -c ${NCU_COUNT:-3} -f -o ${O}_prof $SHORT > ${O}_ncu.log 2>&1
      echo "rc=$?"; tail -n 5 ${O}_ncu.log ;;
    sanitize)
      timeout 2400 compute-sanitizer --tool $arg --error-exitcode 9 --print-limit 20 python -m pytest tests -m gpu -x -q -k "$SUBSET" > ${O}_sanitize_$arg.log 2>&1
      echo "rc=$?"; tail -n 25 ${O}_sanitize_$arg.log ;;
    *) echo "unknown step $step" ;;
  esac
done
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv > ${O}_smi.csv 2>&1
echo "=== done ($(date +%T))"
