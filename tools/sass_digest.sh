#!/bin/bash
# SASS evidence per kernel of libnetsdb_b200.so: tcgen05 (UTC*MMA), TMEM loads (LDTM), TMA (UTMALDG / UBLKCP), legacy tensor ops
cd "$(dirname "$0")/.."
cuobjdump -sass netsdb_b200/libnetsdb_b200.so | awk '
/Function :/ { name=$3; gsub(/^_Z[0-9N]*/,"",name) }
/UTCHMMA/ {u[name]++} /UTCHMMA.*2CTA/ {u2[name]++} /LDTM/ {l[name]++} /UTMALDG/ {t[name]++} /UTMALDG.*MULTICAST/ {tm[name]++}
/UBLKCP/ {b[name]++} /DMMA/ {d[name]++} /HMMA/ {h[name]++} /LDG.*256/ {g[name]++}
END { printf "%-90s %8s %6s %5s %8s %6s %7s %5s %8s\n","kernel (mangled tail)","UTCHMMA",".2CTA","LDTM","UTMALDG",".MCAST","UBLKCP","DMMA","LDG.256";
      for (k in u) printf "%-90s %8d %6d %5d %8d %6d %7d %5d %8d\n", substr(k,1,90), u[k], u2[k], l[k], t[k], tm[k], b[k], d[k], g[k];
      for (k in d) if (!(k in u)) printf "%-90s %8d %6d %5d %8d %6d %7d %5d %8d\n", substr(k,1,90), 0,0,0,0,0,0,d[k],g[k] }' | sort
