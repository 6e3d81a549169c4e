#!/bin/bash
# Round-2 run L: the balanced kernel of mid-size Worlds -- bit-identity against step_kernel first, then a sweep of
# (warps, T) variants and slab lengths against the shipped plans.
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "balanced" > $OUT/r02l_balanced_tests.log 2>&1; echo "pytest rc=$?" >> $OUT/r02l_balanced_tests.log; tail -5 $OUT/r02l_balanced_tests.log
SIZES=${SIZES:-"2000 4000 4736 6000 8000 10000 12000 14208 16384 18944 20000 24000 28000 32000"}
B=nbody_b200/bin/nbody-bench
{
echo "== shipped"
for n in $SIZES; do timeout 60 $B --gpu --fresh-seed --n $n | tail -1; done
for v in ${VARIANTS:-b16x1 b16x2 b16x3 b16x4}; do
  for s in ${SLABS:-0 128 256 512}; do
    echo "== $v slab $s"
    for n in $SIZES; do NBODY_VARIANT=$v NBODY_SLAB=$s timeout 60 $B --gpu --fresh-seed --n $n 2>&1 | tail -1; done
  done
done
} > $OUT/r02l_balanced_sweep.txt 2>&1
python3 - <<'P'
rows={}; cur=None; names=[]
for l in open('gpurun_out/r02l_balanced_sweep.txt'):
    if l.startswith('=='):
        cur=l[3:].strip(); names.append(cur); continue
    f=l.split()
    if len(f)>=3 and f[0].isdigit(): rows.setdefault(int(f[0]),{})[cur]=float(f[2])
print('N'.rjust(6),' '.join(n.replace(' slab ','/').rjust(9) for n in names))
for n in sorted(rows): print(str(n).rjust(6),' '.join(('%9.1f'%rows[n].get(k,float('nan'))) for k in names))
P
