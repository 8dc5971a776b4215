#!/bin/bash
o=gpurun_out/r02_sweep_slab_fma.txt
bash tools/ab_variant.sh $o _variants/fma/libvortex-b200.so > /dev/null 2>&1
for wl in primary_4k_1m cfg1_1080p_primary+shadow_100k; do
  python tools/dump_hits.py - $wl /tmp/a.npy >> $o 2>&1
  python tools/dump_hits.py _variants/fma/libvortex-b200.so $wl /tmp/b.npy >> $o 2>&1
  python - >> $o <<PY
import numpy as np
a, b = np.load('/tmp/a.npy'), np.load('/tmp/b.npy')
ia, ib = np.load('/tmp/a.npy.img.npy'), np.load('/tmp/b.npy.img.npy')
same = a.tobytes() == b.tobytes()
d = np.flatnonzero((a['triIdx'] != b['triIdx']) | (a['dist'] != b['dist']))
print('$wl: exact slab vs FMA slab, whole frame:', len(a), 'primary rays, records that differ:', len(d), '; of those with different t:', int((a['dist'][d] != b['dist'][d]).sum()), '; pixels that differ:', int((ia != ib).sum()))
for k in d[:8]:
    print('   ray', k, 'exact', a[k], 'fma', b[k])
PY
done
cat $o
