#!/bin/bash
# tools/probe_fetch.sh: DRAM bytes per random gather (k_random_gather, 8 GiB table) at the device's default L2 fetch granularity
# and with cudaLimitMaxL2FetchGranularity = 32 / 64 / 128, from ncu's dram__bytes_read.sum -> gpurun_out/l2_fetch.txt
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
out=gpurun_out/l2_fetch.txt
: > $out
for g in 0 32 64 128; do
  for chunk in 32 64 128; do
    DSB_L2_FETCH=$g DSB_VERBOSE=1 ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum,lts__t_sectors_srcunit_tex_op_read.sum --clock-control none -k regex:k_random_gather -s 1 -c 1 --csv \
      python -c "
import sys, ctypes as C
sys.path.insert(0, 'desamba-so_b200')
import pydsb
cuda = pydsb.load_cuda_layer()
t = C.c_float(0)
rc = cuda.dsb_random_access_probe(0, C.c_uint64(8 << 30), C.c_uint32($chunk), C.c_uint64(1 << 26), 2, C.byref(t))
print('PROBE fetch=$g chunk=$chunk rc=%d ms=%.3f gathers/s=%.2fG' % (rc, t.value, (1 << 26) / t.value / 1e6))
" > gpurun_out/l2f.tmp 2>&1
    grep -h "PROBE\|L2 fetch\|dram__bytes_read\|gpu__time\|lts__t_sectors" gpurun_out/l2f.tmp | sed 's/^.*k_random_gather[^,]*,//' | cut -c1-200 >> $out
  done
done
cat $out
