// Lane-group pairing, eight program lanes of two threads each per pairing check: the smallest batches.
#define UC_PROGRAM_INC "pairing_ucode_g8.inc"
#define UCS_KERNEL k_pair_ucode_s8
#define UCS_KERNEL_DYN k_pair_ucode_s8_dyn
#define UCS_LAUNCH launch_pair_ucode_s8
#include "ucode_split.cuh"
