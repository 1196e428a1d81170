# GPU box: full parity suite, both bench arms, the all-configuration and pairs-per-call benches, then ONE ncu pass over
# the bench command (after the same command has exited 0 without ncu):
#   MODE=list (default)  launch list with gpu__time_duration.sum           -> gpurun_out/launches.csv
#   MODE=full            --set full capture of the three matcher kernels   -> summarised on the box by tools/ncu_summary.py
#                        into gpurun_out/ncu_full.csv + gpurun_out/ncu_traffic.json (the .ncu-rep is deleted: gpurun_out/
#                        comes back only if it stays under 64 MiB)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
python bench.py --impl reference --steps 3 --warmup 1 2>gpurun_out/bench_ref.err | tail -1 | tee gpurun_out/bench_ref.json
python bench.py 2>gpurun_out/bench.err | tail -1 | tee gpurun_out/bench.json
python tools/bench_configs.py > gpurun_out/configs.jsonl 2>/dev/null
python tools/batch_curve.py > gpurun_out/batch_curve.jsonl 2>/dev/null
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu --no-video --no-parity"
$BENCH > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; exit 1; }
case "${MODE:-list}" in
  list) ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $BENCH > gpurun_out/ncu_bench.log 2>&1 ;;
  full) ncu --set full --clock-control none -k regex:"k_cost|k_rows|k_vert" --launch-skip 9 -c 3 -o gpurun_out/prof_full -f $BENCH > gpurun_out/ncu_bench.log 2>&1
        python tools/ncu_summary.py gpurun_out/prof_full.ncu-rep 150 gpurun_out/ncu_full.csv --traffic > gpurun_out/ncu_sum.log 2>&1
        cp profiles/ncu_traffic.json gpurun_out/ncu_traffic.json; rm -f gpurun_out/prof_full.ncu-rep ;;
esac
echo "ncu rc=$?"
