#!/bin/bash
# experiment harness: time the lane-group pairing kernel of each library variant under variants/*/ over batch sizes
for v in "$@"; do
  lib=/root/repo/variants/$v/libdrynx_gpu.so
  [ "$v" = default ] && lib=/root/repo/drynx_b200/libdrynx_gpu.so
  echo "== $v"
  DRYNX_LIB=$lib DRX_UC_LANES=${LANES:-4} DRX_RP_PATH=2 timeout 300 python tools/time_verify.py ${SIZES:-100 296 394 500 789} 2>&1 | tr "|" "\n" | grep -o "n=[0-9]*\|pair_ucode=[0-9.]*" | paste - - | tr "\n" ";"
  echo
done
