#!/bin/bash
# SASS evidence that the conv engines are tcgen05 / TMEM / TMA code (no GPU needed):
#   scripts/sass_counts.sh > profiles/sass_counts_rNN.json
SO=vcm_ts_b200/libvcmts_b200.so
T=$(mktemp)
cuobjdump -sass "$SO" > "$T" 2>/dev/null
printf '{"library": "%s", "command": "cuobjdump -sass | grep -c <mnemonic>",\n' "$SO"
for m in UTCHMMA UTCBAR LDTM UTMALDG UTMAPF SYNCS HMMA IMMA FFMA; do
  printf ' "%s": %s,\n' "$m" "$(grep -cw "$m" "$T")"
done
printf ' "kernels_with_UTCHMMA": [%s]\n}\n' "$(awk '/Function :/ {f=$3} /UTCHMMA/ {k[f]++} END {n=0; for (x in k) {printf "%s\"%s: %d\"", (n++ ? ", " : ""), x, k[x]}}' "$T" | c++filt -p 2>/dev/null | sed 's/_Z[A-Za-z0-9_]*/&/g')"
rm -f "$T"
