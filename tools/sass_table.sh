#!/bin/bash
# SASS mnemonic counts per object of libvoxcuda.so -> markdown table on stdout (profiles/r0N_sass_table.md)
cd "$(dirname "$0")/../voxtral.c_b200/build" || exit 1
echo "| object | UTCHMMA (tcgen05.mma) | LDTM (tcgen05.ld) | UTMALDG (TMA tensor) | UBLKCP (TMA bulk) | SYNCS (mbarrier) | HMMA (mma.sync) | LDSM | FFMA2 |"
echo "|---|---:|---:|---:|---:|---:|---:|---:|---:|"
for o in *.o; do
    s=$(cuobjdump -sass "$o" 2>/dev/null)
    row="| \`$o\`"
    for m in UTCHMMA LDTM UTMALDG UBLKCP SYNCS HMMA LDSM FFMA2; do
        if [ "$m" = HMMA ]; then n=$(grep -c "[^C]HMMA" <<<"$s"); else n=$(grep -c "$m" <<<"$s"); fi
        row="$row | $n"
    done
    echo "$row |"
done
