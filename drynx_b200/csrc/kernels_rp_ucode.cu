// Lane-group pairing, four lanes per pairing check (see ucode_body.cuh). Also hosts the per-item prep kernel.
#define UC_PROGRAM_INC "pairing_ucode_g4.inc"
#define UC_KERNEL k_pair_ucode_g4
#define UC_LAUNCH launch_pair_ucode_g4
#define UC_WITH_PREP 1
#define UC_WITH_DYN 1
#define UC_KERNEL_DYN k_pair_ucode_g4_dyn
#include "ucode_body.cuh"
