#!/usr/bin/env bash
# oracle/build_ref.sh - compile the UNMODIFIED reference BLAS path from
# /root/reference (sources stay where they are; nothing is copied into the
# repo) into oracle/_ref/:
#     libvoxref.so   every reference .c except main.c, + oracle/ref_shim.c
#     voxtral_ref    the reference CLI (main.c)
#     lib*openblas*  a copy of the OpenBLAS shared object it links (so the
#                    directory travels to the GPU box as one unit)
#
# Flags follow `make blas` (/root/reference/Makefile:5,53) with ONE change:
# -march=x86-64-v3 instead of -march=native, because the prebuilt objects must
# also run on the GPU box's host CPU (which is not this container's CPU).
# Test infrastructure only; the product never loads anything from here.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${VOX_REFERENCE_DIR:-/root/reference}"
OUT="$HERE/_ref"
INC="$HERE/../include"

if [ ! -f "$REF/voxtral.c" ]; then
    echo "build_ref: $REF not present - keeping prebuilt $OUT (if any)"; exit 0
fi
mkdir -p "$OUT"

SP="$(python3 - <<'EOF'
import site, sys
print(site.getsitepackages()[0])
EOF
)"
BLAS_DEF=""; BLAS_LINK=""
SCIPY_BLAS="$(ls "$SP"/scipy.libs/libscipy_openblas-*.so 2>/dev/null | head -1 || true)"
if [ -n "$SCIPY_BLAS" ]; then
    cp -f "$SCIPY_BLAS" "$OUT/"
    # its private deps (libgfortran / libquadmath) are resolved via RUNPATH=$ORIGIN
    for dep in $(ldd "$SCIPY_BLAS" | awk '/scipy.libs/ {print $3}'); do cp -f "$dep" "$OUT/"; done
    BLAS_DEF="-DVOX_REF_SCIPY_OPENBLAS"
    BLAS_LINK="$OUT/$(basename "$SCIPY_BLAS")"
    echo "build_ref: BLAS = $(basename "$SCIPY_BLAS") (scipy wheel OpenBLAS)" | tee "$OUT/BLAS.txt"
else
    echo "build_ref: BLAS = oracle/sgemm_fallback.c (no OpenBLAS found)" | tee "$OUT/BLAS.txt"
fi

CFLAGS="-O3 -march=x86-64-v3 -ffast-math -fopenmp -DUSE_BLAS -DUSE_OPENBLAS $BLAS_DEF -I$HERE -I$REF -I$INC -fPIC -w"
SRCS="voxtral.c voxtral_kernels.c voxtral_audio.c voxtral_encoder.c voxtral_decoder.c voxtral_tokenizer.c voxtral_safetensors.c voxtral_mic_macos.c"
OBJS=""
for s in $SRCS; do
    o="$OUT/${s%.c}.o"
    gcc $CFLAGS -c "$REF/$s" -o "$o"
    OBJS="$OBJS $o"
done
gcc $CFLAGS -c "$HERE/ref_shim.c" -o "$OUT/ref_shim.o"
gcc $CFLAGS -c "$HERE/sgemm_fallback.c" -o "$OUT/sgemm_fallback.o"
gcc $CFLAGS -c "$REF/main.c" -o "$OUT/main.o"

# --wrap: calls the reference's stream path makes into vox_decoder_forward /
# vox_encoder_forward_incremental pass through the recorders of ref_shim.c (ids, margins,
# rows per encoder call); the functions themselves are the reference's, unmodified.
WRAP="-Wl,--wrap=vox_decoder_forward -Wl,--wrap=vox_encoder_forward_incremental"
gcc -shared -fopenmp $WRAP -o "$OUT/libvoxref.so" $OBJS "$OUT/ref_shim.o" "$OUT/sgemm_fallback.o" \
    $BLAS_LINK -lm -Wl,-rpath,'$ORIGIN'
gcc -fopenmp -o "$OUT/voxtral_ref" "$OUT/main.o" $OBJS "$OUT/sgemm_fallback.o" \
    $BLAS_LINK -lm -Wl,-rpath,'$ORIGIN'
# Strict-IEEE variant (no -ffast-math, no vector libm): same sources, same BLAS.
# -ffast-math lets gcc re-associate the DFT-table angle 2*pi*k*n/400 and call
# libmvec's 4-ulp vector cosf/sinf (voxtral_audio.c:531-537), which moves table
# entries by up to ~1e-4; the strict build is what the C source says, evaluated
# in IEEE float order, and is the pin for the 1e-5 mel tolerance.
SCFLAGS="-O2 -march=x86-64-v3 -fno-fast-math -ffp-contract=off -fopenmp -DUSE_BLAS -DUSE_OPENBLAS $BLAS_DEF -I$HERE -I$REF -I$INC -fPIC -w"
SOBJS=""
for s in $SRCS; do
    o="$OUT/strict_${s%.c}.o"
    gcc $SCFLAGS -c "$REF/$s" -o "$o"
    SOBJS="$SOBJS $o"
done
gcc $SCFLAGS -c "$HERE/ref_shim.c" -o "$OUT/strict_ref_shim.o"
gcc $SCFLAGS -c "$HERE/sgemm_fallback.c" -o "$OUT/strict_sgemm_fallback.o"
gcc -shared -fopenmp $WRAP -o "$OUT/libvoxref_strict.so" $SOBJS "$OUT/strict_ref_shim.o" "$OUT/strict_sgemm_fallback.o" \
    $BLAS_LINK -lm -Wl,-rpath,'$ORIGIN'
rm -f "$OUT"/*.o
echo "build_ref: built $OUT/libvoxref.so and $OUT/voxtral_ref"
