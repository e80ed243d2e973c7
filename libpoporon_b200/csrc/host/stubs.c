/* stubs.c — link-compatibility symbols for the parts of the reference API that are outside the
 * RS/BCH hot path (LDPC: reference src/ldpc.c; RNG: src/rng.c).  They report failure. */
#include "pp_internal.h"

poporon_config_t *poporon_ldpc_config_create(size_t block_size, poporon_ldpc_rate_t rate,
                                             poporon_ldpc_matrix_type_t matrix_type, uint32_t column_weight,
                                             bool use_soft_decode, bool use_outer_interleave, bool use_inner_interleave,
                                             uint32_t interleave_depth, uint32_t lifting_factor,
                                             uint32_t max_iterations, const int8_t *soft_llr, size_t soft_llr_size,
                                             uint64_t seed)
{
    (void)block_size, (void)rate, (void)matrix_type, (void)column_weight, (void)use_soft_decode;
    (void)use_outer_interleave, (void)use_inner_interleave, (void)interleave_depth, (void)lifting_factor;
    (void)max_iterations, (void)soft_llr, (void)soft_llr_size, (void)seed;
    poporon_cuda_note_error("LDPC (reference src/ldpc.c) is not part of poporon-b200: RS and BCH only");
    return NULL;
}

poporon_config_t *poporon_config_ldpc_default(size_t block_size, poporon_ldpc_rate_t rate)
{
    (void)block_size, (void)rate;
    poporon_cuda_note_error("LDPC (reference src/ldpc.c) is not part of poporon-b200: RS and BCH only");
    return NULL;
}

poporon_config_t *poporon_config_ldpc_burst_resistant(size_t block_size, poporon_ldpc_rate_t rate)
{
    (void)block_size, (void)rate;
    poporon_cuda_note_error("LDPC (reference src/ldpc.c) is not part of poporon-b200: RS and BCH only");
    return NULL;
}

poporon_rng_t *poporon_rng_create(poporon_rng_type_t type, void *seed, size_t seed_size)
{
    (void)type, (void)seed, (void)seed_size;
    poporon_cuda_note_error("poporon_rng_* (reference src/rng.c) is not part of poporon-b200");
    return NULL;
}

void poporon_rng_destroy(poporon_rng_t *rng)
{
    (void)rng;
}

bool poporon_rng_next(poporon_rng_t *rng, void *dest, size_t size)
{
    (void)rng, (void)dest, (void)size;
    return false;
}
