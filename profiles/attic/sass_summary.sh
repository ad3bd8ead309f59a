#!/bin/bash
# SASS evidence per object of libmdi_b200.so: counts of the mnemonics that show the sm_100a paths
# (bulk copies / mbarriers, 128-bit loads and stores, multimem, fp64, warp reductions, prefetch)
out=${1:-profiles/r02_sass_summary.txt}
echo "# SASS evidence, libmdi_b200.so at $(git rev-parse --short HEAD) (cuobjdump -sass per object; counts of mnemonics)" > $out
for o in 3dgs_monocular_depth_init_b200/build/*.o; do
  echo "## $(basename $o)" >> $out
  cuobjdump -sass $o 2>/dev/null | grep -oE "\b(UBLKCP[.A-Z0-9]*|UTMA[A-Z.0-9]*|SYNCS[.A-Z0-9]*|LDG\.E\.128[.A-Z0-9]*|LDS\.128|STG\.E\.128[.A-Z0-9]*|STG\.E\.64|STG\.E|MULTIMEM[.A-Z0-9]*|RED[.A-Z0-9]*|ATOMS[.A-Z0-9]*|REDUX[.A-Z0-9]*|FMNMX3|CCTL[.A-Z0-9]*|DADD|DMUL|DFMA|VOTE[.A-Z]*|BAR\.SYNC[.A-Z_]*|MATCH[.A-Z]*)\b" | sort | uniq -c | sort -rn | head -24 >> $out
done
