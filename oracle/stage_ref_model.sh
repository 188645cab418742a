#!/usr/bin/env bash
# TEST INFRASTRUCTURE — not product code.
#
# Stages the reference's own model file (Albatross/rwkv7.py + the CUDA sources it JIT-compiles at import, Albatross/rwkv7.py:47-58)
# byte for byte under the git-ignored oracle/_ref/Albatross/, and pre-builds the torch extension that import needs into
# oracle/_ref/torch_ext/.  /root/reference does not exist on the GPU box; oracle/_ref/ is git-ignored (never in history) but NOT
# gpurun-ignored, so this staged copy travels there exactly like oracle/_ref/libref_wkv7.so does.  It exists so that bench.py's
# `--impl reference` / `cpu_baseline` legs time the REFERENCE'S OWN PyTorch code on the box's host cores (oracle/ref_cpu.py,
# kind "reference") rather than the oracle port.  Nothing in the product path reads it.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REF_ROOT_SRC:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -f "$REF/Albatross/rwkv7.py" ]; then
  echo "stage_ref_model.sh: $REF/Albatross/rwkv7.py not present (expected on the GPU box) — keeping staged files" >&2
  exit 0
fi
mkdir -p "$OUT/Albatross/cuda/sm80"
for f in __init__.py rwkv7.py rwkv_mm_op_triton.py cuda/rwkv7_state_fwd_fp16.cpp cuda/rwkv7_state_fwd_fp16.cu \
         cuda/sm80/rwkv7_state_fwd_fp16.cpp cuda/sm80/rwkv7_state_fwd_fp16.cu; do
  if ! cmp -s "$REF/Albatross/$f" "$OUT/Albatross/$f"; then
    cp -f "$REF/Albatross/$f" "$OUT/Albatross/$f"
    chmod u+w "$OUT/Albatross/$f"
  fi
done
( cd "$REF/Albatross" && sha256sum __init__.py rwkv7.py rwkv_mm_op_triton.py cuda/*.c* cuda/sm80/*.c* ) > "$OUT/Albatross.sha256"
( cd "$OUT/Albatross" && sha256sum -c "$OUT/Albatross.sha256" >/dev/null )
# pre-build the extension the staged module loads at import, with the path spelling the GPU box will use (/root/repo is a symlink
# to the snapshot there), in both variants the module can pick (Albatross/rwkv7.py:40-58): no GPU visible -> the cuda/sm80
# fallback sources; GPU visible -> cuda/ (here: is_available() answered for it, the build itself is the module's own load() call)
PY="${PYTHON:-python}"
cd "$HERE/.."
rm -rf "$OUT/torch_ext"
CUDA_VISIBLE_DEVICES="" REF_ROOT="$OUT" TORCH_CUDA_ARCH_LIST=10.0a "$PY" -c "
from oracle import ref_cpu
R = ref_cpu.import_reference()
print('staged reference imports (no-GPU variant) from', R.__file__)
" 2>&1 | tail -1
CUDA_VISIBLE_DEVICES="" REF_ROOT="$OUT" TORCH_CUDA_ARCH_LIST=10.0a "$PY" -c "
import torch
torch.cuda.is_available = lambda: True
torch.cuda.get_device_capability = lambda *a: (10, 0)
from oracle import ref_cpu
R = ref_cpu.import_reference()
print('staged reference imports (GPU variant) from', R.__file__)
" 2>&1 | tail -1
