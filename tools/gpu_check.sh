# GPU box: full parity suite, both bench arms, then the ncu launch list of the same bench command.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
python bench.py --impl reference --steps 3 --warmup 1 2>gpurun_out/bench_ref.err | tail -1 | tee gpurun_out/bench_ref.json
python bench.py 2>gpurun_out/bench.err | tail -1 | tee gpurun_out/bench.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_bench.log 2>&1; echo "ncu rc=$?"
