#!/bin/bash
# A/B of the programmatic dependent launch of the short mat-vec kernels: LPN line with and without it.
for e in 0 1; do
  if [ $e = 1 ]; then export M4RI_B200_NO_PDL=1; else unset M4RI_B200_NO_PDL; fi
  python bench.py --workload lpn --steps 20 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('NO_PDL=$e matvec_us', round(d['matvec']['ms']*1e3,2), 'hbm_frac', round(d['matvec']['roofline']['frac'],3), 'single_flushed_us', round(d['matvec']['ms_flushed_single_launch']*1e3,2), 'batched8_us', round(d['matvec_batched_8']['ms']*1e3,2), 'batched32_us', round(d['matvec_batched_32']['ms']*1e3,2), '2^22_frac', round(d['matvec_2^22_rows']['roofline']['frac'],3))"
done
