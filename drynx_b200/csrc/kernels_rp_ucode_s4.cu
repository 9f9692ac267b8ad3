// Lane-group pairing, four program lanes of two threads each per pairing check (see ucode_split.cuh).
#define UC_PROGRAM_INC "pairing_ucode_g4.inc"
#define UCS_KERNEL k_pair_ucode_s4
#define UCS_KERNEL_DYN k_pair_ucode_s4_dyn
#define UCS_LAUNCH launch_pair_ucode_s4
#include "ucode_split.cuh"
