# GPU box: full parity suite, both bench arms, then ONE ncu pass over the same bench command:
#   MODE=list (default)  launch list with gpu__time_duration.sum           -> gpurun_out/launches.csv
#   MODE=traffic         DRAM bytes / instructions of the matcher kernels  -> gpurun_out/traffic.csv
#   MODE=full            --set full capture of the three matcher kernels   -> gpurun_out/prof_full.ncu-rep
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
python bench.py --impl reference --steps 3 --warmup 1 2>gpurun_out/bench_ref.err | tail -1 | tee gpurun_out/bench_ref.json
python bench.py 2>gpurun_out/bench.err | tail -1 | tee gpurun_out/bench.json
case "${MODE:-list}" in
  list) ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
          python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_bench.log 2>&1 ;;
  traffic) ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,smsp__inst_executed.sum,lts__t_sector_hit_rate.pct \
          --clock-control none -k regex:"k_cost|k_rows|k_vert" --launch-skip 9 -c 3 --csv --log-file gpurun_out/traffic.csv \
          python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_bench.log 2>&1 ;;
  full) ncu --set full --clock-control none --import-source on -k regex:"k_cost|k_rows|k_vert" --launch-skip 9 -c 3 \
          -o gpurun_out/prof_full -f python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_bench.log 2>&1 ;;
esac
echo "ncu rc=$?"
