#!/bin/bash
# usage: tools/sass_stats.sh lib.so [regex]  -> per-kernel SASS size and uniform-datapath evidence
cuobjdump -sass "$1" 2>/dev/null | awk -v pat="${2:-.}" '
/Function : /{f=$3}
/^ +\/\*[0-9a-f]+\*\/ /{n[f]++; if ($0 ~ / LDCU/) u[f]++; if ($0 ~ / LDC[ .]/) l[f]++; if ($0 ~ /BSSY/) b[f]++; if ($0 ~ /BRA\.U/) bu[f]++; else if ($0 ~ / BRA /) br[f]++; if ($0 ~ / (LDL|STL)/) loc[f]++}
END{for (f in n) if (f ~ pat) printf "%6d ins  LDC %4d  LDCU %4d  BSSY %4d  BRA.U %4d  BRA %4d  LDL/STL %4d  %s\n", n[f], l[f], u[f], b[f], bu[f], br[f], loc[f], substr(f,1,60)}' | sort -k16
