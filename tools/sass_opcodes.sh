#!/bin/bash
# histogram of the SASS opcodes that matter (FP64 tensor pipe, bulk copies / mbarriers, FP64 pipe, local-memory spills) per kernel
# of the shipped library: tools/sass_opcodes.sh > profiles/sass_opcodes.txt
LIB=1d-burgers-equation-roms_b200/libburgers_b200.so
cuobjdump -sass $LIB | awk '
/Function :/ { f=$3 }
/DMMA/ { dmma[f]++ } /UBLKCP/ { blk[f]++ } /SYNCS/ { syncs[f]++ } /DFMA/ { dfma[f]++ } /DMUL|DADD/ { dadd[f]++ }
/MUFU.RCP64H/ { rcp[f]++ } /SHFL/ { shfl[f]++ } /LDL/ { ldl[f]++ } /STL/ { stl[f]++ } /BAR.SYNC|BAR.ARV/ { bar[f]++ } /BRA.DIV/ { div[f]++ }
/REDUX/ { redux[f]++ } /LDG/ { ldg[f]++ } /LDS/ { lds[f]++ }
{ if (f != "") seen[f]=1 }
END { printf "%-64s %6s %6s %6s %6s %6s %6s %6s %6s %6s %5s %5s %5s %5s\n", "kernel (c++filt)", "DMMA", "UBLKCP", "SYNCS", "DFMA", "DMUL+DADD", "RCP64H", "SHFL", "LDS", "LDG", "BAR", "LDL", "STL", "BRA.DIV";
  for (k in seen) printf "%-64s %6d %6d %6d %6d %6d %6d %6d %6d %6d %5d %5d %5d %5d\n", k, dmma[k], blk[k], syncs[k], dfma[k], dadd[k], rcp[k], shfl[k], lds[k], ldg[k], bar[k], ldl[k], stl[k], div[k] }' | (read hdr; echo "$hdr"; sort) | while read -r name rest; do if [ "$name" = "kernel" ]; then echo "$name $rest"; else printf "%-64s %s\n" "$(echo $name | c++filt | sed 's/(bx::.*//; s/^void //; s/bx:://' | cut -c1-64)" "$rest"; fi; done
