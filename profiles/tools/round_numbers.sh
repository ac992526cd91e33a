#!/bin/bash
# The numbers DESIGN.md / BASELINE.md quote, from one box: default bench line (with the end-to-end legs),
# the reference arm, the other configs and hash modes (device-resident scan only), small records.
o=gpurun_out
python bench.py > $o/bench_n1.json 2> $o/bench_n1.err || echo "bench failed"
python bench.py --impl reference > $o/bench_ref.json 2> $o/bench_ref.err || echo "reference arm failed"
: > $o/other_configs.txt
for w in cfg2_10k_templates_4KB_x8_nodrift cfg3_100k_configmaps_64KB_x16_1pct cfg5_10k_secrets_1MB_x4; do
  python bench.py --workload $w --skip-e2e --skip-cpu --steps 30 2>/dev/null | grep '^{' | tail -1 >> $o/other_configs.txt
done
for m in hash64 xxh64 sha256; do
  python bench.py --mode $m --skip-e2e --skip-cpu --steps 10 2>/dev/null | grep '^{' | tail -1 >> $o/other_configs.txt
done
rm -f $o/small_final.txt
profiles/tools/sweep_small.sh $o/small_final.txt
python - <<'P'
import json
for l in open('gpurun_out/other_configs.txt'):
    d = json.loads(l); r = d['roofline']
    print(d['config']['workload'], d['config']['mode'], round(d['ms_per_step'], 4), 'ms', f"{d['value']:.3g}", 'dec/s', round(d['gb_per_s']), 'GB/s whole;', r['kernel'], round(r['achieved']), round(r['frac'], 3), r['phases_ms'])
P
cat $o/small_final.txt
