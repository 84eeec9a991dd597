#!/bin/bash
# The non-headline shapes of SURVEY.md 8(d) (parity-test cases, not bench lines): one line per workload / corpus.
# usage (on the GPU box): bash tools/other_workloads.sh > gpurun_out/r1_other_workloads.txt
for spec in "c1 words" "c3 words" "c3 az09" "c3 rand62" "c4 words" "c5 words" "c5 az09" "c5 rand62" "c2 az09" "c2 rand62"; do
  set -- $spec
  timeout 600 python bench.py --workload $1 --corpus $2 --steps 3 --warmup 3 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); o=d['ops']
print('%-3s %-7s payload %6.1f MB  step %8.3f ms = %6.1f GB/s | crc %.3f compress %.3f pack %.3f decompress %.3f ms | e2e %.2f GB/s (3 calls %.2f) | ratio %.3f' % (
  '$1','$2', d['config']['payload_bytes_per_gpu']/1e6, d['ms_per_step'], d['value'], o['crc32c']['ms_per_step'], o['compress']['ms_per_step'], o['pack']['ms_per_step'], o['decompress']['ms_per_step'], d['e2e']['value'], d['e2e']['separate_calls_value'], d['compression_ratio']))"
done
