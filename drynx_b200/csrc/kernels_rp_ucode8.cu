// Lane-group pairing, eight lanes per pairing check: half the product phases per lane, for the smallest batches.
#define UC_PROGRAM_INC "pairing_ucode_g8.inc"
#define UC_KERNEL k_pair_ucode_g8
#define UC_LAUNCH launch_pair_ucode_g8
#include "ucode_body.cuh"
