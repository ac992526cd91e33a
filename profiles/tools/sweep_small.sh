#!/bin/bash
# Small-record scans: 1M records x 8 shards (shard copies in shuffled order), per record length, with the
# experiment knobs given as VAR=value arguments after the output file (e.g. NCD_LINE_CACHE=0 NCD_SS_LPO=8).
out=${1:-gpurun_out/small_sweep.txt}; shift
for kv in "$@"; do export "$kv"; done
echo "# knobs: $*" >> $out
for len in ${LENS:-64 128 256 512 1024}; do
  line=$(NCD_BENCH_LEN=$len,$len python bench.py --skip-e2e --skip-cpu --steps 20 --warmup 3 2>/dev/null | grep '^{' | tail -1)
  echo "$len $(echo "$line" | python -c 'import json,sys; d=json.loads(sys.stdin.read()); r=d["roofline"]; print(round(d["ms_per_step"],4), r["kernel"], round(r["kernel_ms"],4), round(d["gb_per_s"]), r["phases_ms"])')" >> $out
done
