#!/bin/bash
# Experiment: does the inlined, fully unrolled SHA3-512 still give wrong challenges inside the proof-header kernels?
# Builds a variant library into variants/keccak_inline/ (git-ignored) and runs the creation tests against it.
set -e
cd "$(dirname "$0")/.."
mkdir -p variants/keccak_inline
DRX_OBJDIR=variants/keccak_inline DRYNX_LIB=$PWD/variants/keccak_inline/libdrynx_gpu.so DRX_NVCC_EXTRA="-DDRX_KECCAK_INLINE=1" \
  python -m drynx_b200.build gpu
echo "variant built: variants/keccak_inline/libdrynx_gpu.so"
if nvidia-smi -L >/dev/null 2>&1; then
  DRYNX_LIB=$PWD/variants/keccak_inline/libdrynx_gpu.so python -m pytest tests/test_gpu_rangeproof.py -m gpu -q -x -k "create_bytes or deterministic or creation_kernels" 2>&1 | tail -15
fi
