#!/usr/bin/env bash
# Builds the UNMODIFIED reference extension for sm_100a into baseline/_ref (git-ignored,
# travels to the GPU box with gpurun).  /root/reference is read-only, so the build runs in
# a scratch copy under /tmp; only the Python package dir (+ the built _C*.so) is kept.
# Used by: tests/golden/gen_golden.py, bench.py --impl reference, tools/ref_probe.py.
set -euo pipefail
REF=${1:-/root/reference}
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
OUT="$ROOT/baseline/_ref"
if [ ! -d "$REF" ]; then echo "no reference tree at $REF; keeping existing $OUT"; exit 0; fi
if ls "$OUT"/diff_gaussian_sampling/_C*.so >/dev/null 2>&1; then echo "reference already built"; exit 0; fi
TMP=$(mktemp -d /tmp/refbuild.XXXXXX)
mkdir -p "$TMP/src"
# copy everything except glm's docs/tests (dead weight; glm itself is an unused include)
tar -C "$REF" --exclude='third_party/glm/doc' --exclude='third_party/glm/test' -cf - . | tar -C "$TMP/src" -xf -
cd "$TMP/src"
export TORCH_CUDA_ARCH_LIST=10.0a MAX_JOBS=${MAX_JOBS:-6}
python setup.py build_ext --inplace > "$TMP/build.log" 2>&1 || { tail -50 "$TMP/build.log"; exit 1; }
mkdir -p "$OUT"
rm -rf "$OUT/diff_gaussian_sampling"
cp -r "$TMP/src/diff_gaussian_sampling" "$OUT/"
ls -la "$OUT/diff_gaussian_sampling"
rm -rf "$TMP"
