#!/bin/bash
# Blackwell-native evidence: SASS mnemonic counts per kernel of the built library
# (tcgen05.mma -> UTC*MMA, tcgen05.ld -> LDTM, TMA -> UTMALDG, tcgen05.commit -> UTCBAR).
# Usage: tools/sass_digest.sh > profiles/r2_sass_digest.txt
so=gabril-carla_b200/libgabril_b200.so
echo "# cuobjdump -sass $so ($(cuobjdump -lelf $so | head -1 | tr -d '\n'))"
echo "# kernel | instructions | UTCHMMA | UTMALDG(2D/3D/4D.IM2COL) | LDTM | UTCBAR | FADD2/FMUL2 | STG.256 | HMMA(legacy)"
cuobjdump -sass $so | awk '
/Function :/ { if (name != "") print_row(); name=$3; n=0; mma=0; tma=0; im2=0; ldtm=0; bar=0; f2=0; st=0; hm=0 }
/^[ \t]+\/\*[0-9a-f]+\*\// { n++ }
/UTCHMMA/ {mma++} /UTMALDG/ {tma++} /IM2COL/ {im2++} /LDTM/ {ldtm++} /UTCBAR/ {bar++} /FADD2|FMUL2/ {f2++} /STG.E.ENL2.256/ {st++} / HMMA/ {hm++}
function print_row() { printf "%s | %d | %d | %d (%d im2col) | %d | %d | %d | %d | %d\n", name, n, mma, tma, im2, ldtm, bar, f2, st, hm }
END { print_row() }' | c++filt | sort
