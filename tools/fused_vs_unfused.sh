for e in 0 1; do
  if [ $e = 1 ]; then export M4RI_B200_NO_FUSED_STRASSEN=1; fi
  python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('nofused=$e ms/step', round(d['ms_per_step'],3), 'kernel_ms', round(r['kernel_ms'],4), 'share', round(r['kernel_share_of_step'],4), 'launches', d['gpu_launches'], d['step_ms_rank0'])"
done
