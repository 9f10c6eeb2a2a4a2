#!/usr/bin/env bash
# A/B of kernel variants on the GPU box: tools/gpu_ab.sh <tag> "<ENV=.. ENV=..>" ...  -> gpurun_out/ab_<tag>.log
# each variant runs bench.py (C3, N=1, device-resident only) and prints value + the stage table
set -u
mkdir -p gpurun_out
tag=$1; shift
out=gpurun_out/ab_${tag}.log
: > $out
for envs in "$@"; do
  echo "=== $envs" >> $out
  env $envs python bench.py --steps ${STEPS:-10} --warmup 3 --no-cpu-baseline --no-ref-cuda --no-e2e ${BENCH_ARGS:-} 2>>$out | python -c "
import sys, json
for line in sys.stdin:
    line=line.strip()
    if not line.startswith('{'): continue
    d=json.loads(line)
    print('value', round(d['value'],3), 'lean', d.get('value_without_unused_geometry_maps'), 'fwd_ms', round(d['fwd_ms'],3), 'clk', d['clocks']['sm_mhz'], d['clocks']['reasons'])
    for k,v in d['stages'].items(): print('  %-14s %8.3f ms  %s GB/s' % (k, v['ms_per_step'], None if v['alg_GBps'] is None else round(v['alg_GBps'])))
" >> $out
done
cat $out
