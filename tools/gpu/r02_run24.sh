#!/bin/bash
# Round 2, GPU call: evict-last stores for the voxel lists -- parity, speed, DRAM bytes (ncu) against plain stores
set -u
O=gpurun_out/${RUN_TAG:-r02aa}
mkdir -p $O
python - <<'PY' > $O/l2_attrs.txt 2>&1
import ctypes
rt = ctypes.CDLL("libcudart.so.12")
for name, attr in (("cudaDevAttrMaxPersistingL2CacheSize", 108), ("cudaDevAttrMaxAccessPolicyWindowSize", 109), ("cudaDevAttrL2CacheSize", 38)):
    v = ctypes.c_int(0)
    rc = rt.cudaDeviceGetAttribute(ctypes.byref(v), attr, 0)
    print(name, rc, v.value)
PY
cat $O/l2_attrs.txt
( timeout 900 python -m pytest tests/test_search.py tests/test_gpu_parity.py -m gpu -q -x ) > $O/pytest_parity.txt 2>&1
tail -3 $O/pytest_parity.txt
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --views-per-step 10000"
for v in keep keep_extra0 nokeep; do
  export ASSCPP_LIB=$PWD/asscpp_b200/lib/libasscpp_b200.so; unset ASSCPP_L2_PERSIST_EXTRA_MB
  if [ $v = nokeep ]; then export ASSCPP_LIB=$PWD/asscpp_b200/lib/variants/lib_nokeep.so; fi
  if [ $v = keep_extra0 ]; then export ASSCPP_L2_PERSIST_EXTRA_MB=0; fi
  timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline > $O/bench_$v.json 2> $O/bench_$v.err
  python -c "import json;d=json.loads(open('$O/bench_$v.json').read().strip().splitlines()[-1]);print('$v',round(d['value']),round(d['e2e']['value']))"
  $CMD > $O/plain_$v.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,smsp__inst_executed.sum --clock-control none -k regex:eval_views_kernel -s 6 -c 1 --csv --log-file $O/ncu_$v.csv $CMD > $O/ncu_$v.log 2>&1
  grep -v "^==" $O/ncu_$v.csv | tail -5 | awk -F'","' '{print $(NF-2), $(NF-1), $NF}'
done
