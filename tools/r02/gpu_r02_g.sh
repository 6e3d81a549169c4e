#!/bin/bash
set -u
OUT=gpurun_out; mkdir -p $OUT
SIZES=1000,2000,3000,4000,4736,5000,6000,8000,9472,10000,11000,12000,13000,14208,15000,16384,18000,18944,20000
{
echo "== step_kernel (NBODY_SMALL_MAX_N=0)"; NBODY_SMALL_MAX_N=0 nbody_b200/bin/nbody-bench --gpu --fresh-seed --n $SIZES
for v in w8x1 w16x1 w32x1 w8x2 w16x2 w32x2 w8x3 w16x3 w8x4 w16x4; do
  echo "== small_world_kernel $v (NBODY_VARIANT)"; NBODY_VARIANT=$v nbody_b200/bin/nbody-bench --gpu --fresh-seed --n $SIZES
done
} > $OUT/r02g_small_world_variants.txt 2>&1
python3 - <<'P'
import re
rows={}; cur=None
for l in open('gpurun_out/r02g_small_world_variants.txt'):
    if l.startswith('=='): cur=l.split()[1] if 'step_kernel' in l else l.split()[2]; continue
    f=l.split()
    if len(f)>=3 and f[0].isdigit(): rows.setdefault(int(f[0]),{})[cur]=float(f[2])
names=['step_kernel','w8x1','w16x1','w32x1','w8x2','w16x2','w32x2','w8x3','w16x3','w8x4','w16x4']
print('N'.rjust(6),' '.join(n.rjust(8) for n in names))
for n in sorted(rows): print(str(n).rjust(6),' '.join(('%8.1f'%rows[n].get(k,float('nan'))) for k in names))
P
