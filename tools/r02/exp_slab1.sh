for s in 0 100000; do echo "== NBODY_SLAB=$s"; NBODY_SLAB=$s nbody_b200/bin/nbody-bench --gpu --fresh-seed --n 500,800,1000,1200,1600,2000,3000,4000,6000; done
for v in t64x4 t128x4; do echo "== NBODY_SLAB=100000 NBODY_VARIANT=$v"; NBODY_VARIANT=$v NBODY_SLAB=100000 nbody_b200/bin/nbody-bench --gpu --fresh-seed --n 500,800,1000,1200,1600,2000,3000,4000,6000; done
