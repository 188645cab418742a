#!/usr/bin/env bash
# TEST INFRASTRUCTURE — not product code.
#
# Builds the *reference's own* CUDA WKV/SPMV kernels for sm_100a from the sources where they lie
# (/root/reference/Albatross/cuda/rwkv7_state_fwd_fp16.cu — never copied into this repo) into
# oracle/_ref/libref_wkv7.so.  Only the .cu is compiled (not the TORCH_LIBRARY .cpp), so loading it
# registers no torch ops; tests call its C++ launchers (cuda_forward_seq / cuda_forward_one /
# cuda_spmv_forward, reference .cu:313-330) through their mangled names with ctypes.
#
# Flags follow the reference's JIT build line (Albatross/rwkv7.py:54-55) plus the sm_100a gencode.
# oracle/_ref/ is git-ignored but NOT gpurun-ignored: the .so travels to the GPU box, where it is the
# kernel-level A/B checker for tests/ (bit-exact comparison) and nothing else.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REF_ROOT:-/root/reference}"
SRC="$REF/Albatross/cuda/rwkv7_state_fwd_fp16.cu"
OUT="$HERE/_ref"
if [ ! -f "$SRC" ]; then
  echo "build_ref.sh: $SRC not present (expected on the GPU box) — keeping prebuilt files" >&2
  exit 0
fi
mkdir -p "$OUT"
PY="${PYTHON:-python}"
TORCH_DIR="$($PY -c 'import torch, os; print(os.path.dirname(torch.__file__))' 2>/dev/null)"
if [ "$OUT/libref_wkv7.so" -nt "$SRC" ] && [ "$OUT/libref_wkv7.so" -nt "$0" ]; then
  exit 0
fi
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo \
  -res-usage --use_fast_math -O3 --extra-device-vectorization -D_N_=64 -Xptxas -O3 \
  -std=c++17 -shared -Xcompiler -fPIC \
  -I"$TORCH_DIR/include" -I"$TORCH_DIR/include/torch/csrc/api/include" \
  -L"$TORCH_DIR/lib" -lc10 -lc10_cuda -Xlinker -rpath -Xlinker "$TORCH_DIR/lib" \
  -o "$OUT/libref_wkv7.so" "$SRC"
echo "built $OUT/libref_wkv7.so"
