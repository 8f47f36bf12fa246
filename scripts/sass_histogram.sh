#!/bin/bash
# SASS opcode evidence per kernel (no GPU needed): tcgen05 MMA (UTC*MMA), TMA loads/stores (UTMALDG/UTMASTG), TMEM loads (LDTM),
# TMEM alloc (UTCATOMSWS / UTCBAR ...) counted per kernel of the built library.   bash scripts/sass_histogram.sh > profiles/<tag>_sass_opcodes.txt
LIB=${1:-dincae.jl_b200/libdincae_b200.so}
cuobjdump -sass "$LIB" | awk '
  /Function :/ { fn=$3 }
  { while (match($0, /UTC[A-Z]*MMA[.A-Z0-9_]*|UTMALDG[.A-Z0-9_]*|UTMASTG[.A-Z0-9_]*|LDTM[.A-Z0-9_]*|STTM[.A-Z0-9_]*|UTCBAR[.A-Z0-9_]*|UTCATOMSWS[.A-Z0-9_]*|SYNCS[.A-Z0-9_]*/)) {
      op=substr($0, RSTART, RLENGTH); n[fn" "op]++; $0=substr($0, RSTART+RLENGTH) } }
  END { for (k in n) print n[k], k }' | sort -k2,2 -k1,1nr | c++filt | awk '{c=$1; $1=""; printf "%6d %s\n", c, $0}'
