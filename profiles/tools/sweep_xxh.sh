#!/bin/bash
# k_xxh64 at cfg 4 (1M records of 1-16 KiB x 9 arenas) per staging variant: "WIN NBUF STAGE" triples
# (window bytes; row sets per warp; 0 = one bulk copy per lane, 1 = cooperative 16-byte cp.async).
out=${1:-gpurun_out/xxh_sweep.txt}
: > $out
while read win nbuf stage; do
  [ -z "$win" ] && continue
  NCD_NVCC_EXTRA="-DNCD_XXH_WIN=$win -DNCD_XXH_NBUF=$nbuf -DNCD_XXH_STAGE=$stage" python -c "
from nexus_configuration_controller_b200 import build
build.build_native(force=True)" >/dev/null 2>&1
  python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "xxh or hash" 2>&1 | tail -1 >> $out
  line=$(python bench.py --mode xxh64 --skip-e2e --skip-cpu --steps 10 --warmup 3 2>/dev/null | grep '^{' | tail -1)
  echo "win=$win nbuf=$nbuf stage=$stage $(echo "$line" | python -c 'import json,sys; d=json.loads(sys.stdin.read()); r=d["roofline"]; print(round(d["ms_per_step"],3), r["kernel"], round(r["kernel_ms"],3), round(r["achieved"]), round(r["frac"],3))')" >> $out
done <<< "${VARIANTS:-512 1 0
512 1 1
256 2 1
256 1 1
512 2 1}"
python -c "
from nexus_configuration_controller_b200 import build
build.build_native(force=True)" >/dev/null 2>&1
cat $out
