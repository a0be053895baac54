#!/bin/bash
# Per-kernel counts of the Blackwell-specific SASS mnemonics in the shipped library: UTCIMMA(.2CTA) = tcgen05.mma.kind::i8,
# UTMALDG = TMA tensor loads, UBLKCP = cp.async.bulk, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, SYNCS = mbarrier,
# UTCATOMSWS = TMEM allocation.   tools/sass_listing.sh > profiles/sass_tcgen05_rNN.txt
cd "$(dirname "$0")/.."
cuobjdump -sass fricc_b200/libfricc_b200.so 2>/dev/null \
 | grep -E "Function :|UTCIMMA|UTCBAR|UTMALDG|LDTM|UTCCP|UTMACCTL|UBLKCP|SYNCS|UTCATOM|UTCHMMA|UTCQMMA|RED.E.ADD.F64" \
 | awk '/Function :/{fn=$3; next} {split($0,a,";"); ins=a[1]; sub(/^ *\/\*[0-9a-f]+\*\/ */,"",ins); n=split(ins,w," "); key=w[1]; if (w[1] ~ /^@/) key=w[2]; cnt[fn" "key]++} END{for(k in cnt) print k, cnt[k]}' \
 | sort | c++filt | awk '{c=$NF; $NF=""; printf "%6d  %s\n", c, $0}'
