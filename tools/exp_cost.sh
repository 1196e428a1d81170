cd depthestimation_b200/csrc
for MB in 3 4 6; do
  /usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC --fmad=false -DCOST_MINB=$MB -c cost.cu -o cost.o 2>/dev/null
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libb200stereo.so api.o cost.o paths.o wta.o imgops.o speckle.o rows.o
  for XT in 16 32; do for RS in 32 64; do
    echo -n "MINB=$MB XT=$XT RS=$RS: "
    (cd ../..; B200S_COST_XT=$XT B200S_COST_RS=$RS python bench.py --steps 3 --warmup 3 --no-cpu 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('cost ms', round(d['stage_ms_per_step']['cost'],3), 'fps', round(d['value']))")
  done; done
done
