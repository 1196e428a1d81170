# row-kernel tuning on the GPU box: min-blocks-per-SM (register cap) x batch
cd depthestimation_b200/csrc
OBJS="api.o cost.o paths.o wta.o imgops.o speckle.o rows.o vert.o"
for MB in ${MBS:-4 5 6}; do
  /usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC --fmad=false -DROWS_MINB=$MB $EXTRA -Xptxas -v -c rows.cu -o rows.o 2> /tmp/rows.log
  grep -A1 "k_rowsILi16ELi8ELb0" /tmp/rows.log | grep -E "registers|spill" | tr '\n' ' '
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libb200stereo.so $OBJS
  for BT in ${BATCHES:-30}; do
    echo -n "MINB=$MB B=$BT: "
    (cd ../..; python bench.py --steps 3 --warmup 3 --no-cpu --batch $BT 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); s=d['stage_ms_per_step']; n=d['config']['pairs_per_step_per_gpu']; print('per-frame us: wta', round(1e3*s['wta']/n,1), 'paths', round(1e3*s['paths']/n,1), 'cost', round(1e3*s['cost']/n,1), 'fps', round(d['value']))")
  done
done
