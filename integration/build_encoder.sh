#!/bin/bash
# Builds SvtAv1EncApp twice from COPIES of the reference tree (C-only: no nasm on this image), out of tree:
#   integration/_build/SvtAv1EncApp_stock   unmodified
#   integration/_build/SvtAv1EncApp_gpu     with the B200 ME binding (integration/apply_patch.py)
# and writes the binding as a unified diff to integration/encoder_me_gpu.patch. The binaries are git-ignored and travel
# to the GPU box with the snapshot, like oracle/_ref/libme_ref.so. /root/reference itself is never written to.
set -e
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
WORK=${WORK:-/tmp/svt_enc_build}
OUT=$HERE/_build
mkdir -p "$WORK" "$OUT"
for v in stock gpu; do
    rm -rf "$WORK/src_$v"
    mkdir -p "$WORK/src_$v"
    (cd "$REF" && tar --exclude=.git --exclude=Bin -cf - .) | tar -xf - -C "$WORK/src_$v"
done
python "$HERE/apply_patch.py" "$WORK/src_gpu"
# (the files' timestamps are cut from the diff headers: the patch changes only when its content does)
(cd "$WORK" && diff -urN src_stock/Source src_gpu/Source | grep -v '^Only in' | sed -E 's/^(--- |\+\+\+ )([^\t]+)\t.*/\1\2/' > "$HERE/encoder_me_gpu.patch" || true)
for v in stock gpu; do
    mkdir -p "$WORK/build_$v"
    (cd "$WORK/build_$v" && cmake "$WORK/src_$v" -G Ninja -DCMAKE_BUILD_TYPE=Release -DCOMPILE_C_ONLY=ON -DBUILD_TESTING=OFF \
        -DBUILD_SHARED_LIBS=OFF -DSVT_AV1_LTO=OFF -DEXCLUDE_HASH=ON > cmake.log 2>&1 && ninja > ninja.log 2>&1) || { tail -30 "$WORK/build_$v/ninja.log"; exit 1; }
    cp "$WORK/src_$v/Bin/Release/SvtAv1EncApp" "$OUT/.tmp_$v" && mv -f "$OUT/.tmp_$v" "$OUT/SvtAv1EncApp_$v"
done
ls -la "$OUT"
