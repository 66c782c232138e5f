#!/bin/bash
# A/B of kernel variants at mapping scale: tools/ab_map.sh name1 name2 ...  ("base" = the product library)
# prints the binning / projection kernels' per-call times of an 8-view, 500 k-Gaussian mapping step and of a 1-view one
for v in "$@"; do
  if [ "$v" = base ]; then lib=gslam_b200/libgsb.so; else lib=build/var_$v/libgsb_$v.so; fi
  for c in 8 1; do
    echo "== $v, $c view(s)"
    GSB_LIB=$lib python tools/map_profile.py $c 500000 2>&1 | grep -E "step:|scatter|segment_sort|project_fwd|tile_scan|graph replay"
  done
done
