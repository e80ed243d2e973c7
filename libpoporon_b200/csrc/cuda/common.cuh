// common.cuh — shared declarations of the libpoporon_cuda translation units.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "poporon_cuda.h"

namespace ppcu {

void set_error(const char *fmt, ...);
int check(cudaError_t e, const char *what);

#define PP_CUDA_TRY(expr)                                   \
    do {                                                    \
        int _rc = ::ppcu::check((expr), #expr);             \
        if (_rc != 0)                                       \
            return _rc;                                     \
    } while (0)

struct DeviceGuard {
    int prev = -1;
    bool ok = false;
    explicit DeviceGuard(int dev)
    {
        if (cudaGetDevice(&prev) != cudaSuccess)
            prev = -1;
        ok = (prev == dev) || (cudaSetDevice(dev) == cudaSuccess);
    }
    ~DeviceGuard()
    {
        if (prev >= 0)
            cudaSetDevice(prev);
    }
};

// Field + code tables as the generic kernels see them (device global memory, copied to shared
// memory by each CTA).  All entries fit a byte for m <= 8: the log of zero is nn <= 255.
struct RsTables {
    uint8_t alpha_to[256];
    uint8_t index_of[256];
    uint8_t genpoly[256]; // index form, nroots+1 entries
};

struct RsCode {
    int m;
    int nn;
    int nroots;
    uint16_t fcr, prim, iprim;
};

// kernel ids of the per-kernel profile (poporon_cuda_rs_profile_read)
enum { PROF_ENCODE = 0, PROF_SCREEN, PROF_COMPACT, PROF_SOLVE, PROF_PARITY, PROF_GENERAL, PROF_GENERIC, PROF_ROWS, PROF_ERASURE, PROF_EE, PROF_KINDS };
constexpr int kProfCap = 8192;

static inline int sm_count_of(int device)
{
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess)
        return 148;
    return n;
}

} // namespace ppcu

struct poporon_cuda_rs;
namespace ppcu {
// RAII event pair around one kernel launch; a no-op unless profiling is on for the context
struct ProfScope {
    poporon_cuda_rs *ctx;
    cudaStream_t stream;
    int slot;
    ProfScope(poporon_cuda_rs *c, int id, cudaStream_t s);
    ~ProfScope();
};
} // namespace ppcu

// Context layouts (opaque to C callers).
struct poporon_cuda_rs {
    int device;
    int sm_count;
    ppcu::RsCode code;
    ppcu::RsTables *d_tables; // device copy
    bool fast255;             // RS(255,223), 0x11D, fcr = 1, prim = 1: bitsliced kernels apply
    bool force_generic;
    uint64_t launches;
    // scratch for the staged decode pipeline (grow-only, owned)
    uint32_t *d_queue;        // dirty-codeword indices
    uint8_t *d_rem;           // per-dirty-codeword remainder / syndromes
    uint32_t *d_counters;     // [0] = dirty count
    size_t queue_cap;
    size_t scratch_words_per_sm; // per-lane scratch words over all the warps of an SM that d_rem has room for
    uint8_t *d_slab;          // rs255_fft: one 255 KB bit-plane slab per SM (allocated on first use)
    uint32_t *d_synplanes;    // rs255_fft encoder: syndrome planes of the chunk in flight (grow-only)
    size_t synplane_cap;      // codewords it holds
    cudaEvent_t busy;         // recorded after the last kernel that uses the scratch above; the next call's stream waits on it
    int correct_ctas_per_sm;  // resident CTAs of the general correction stage (per device, rs255_init)
    int no_fuse;              // PP_RS255_NOFUSE=1: compaction kernel instead of the fused zero-syndrome exit (A/B runs)
    uint32_t ee_min_rows;     // PP_RS255_EE_MIN_ROWS: see rs255_ee_kernel
    int no_ee_kernel;         // PP_RS255_NO_EE_KERNEL=1: erasure mode without its second bitsliced stage (A/B runs)
    int no_erasure_kernel;    // PP_RS255_NO_ERASURE_KERNEL=1: erasure mode without the bitsliced erasure-only kernel (A/B runs)
    int use_lfsr;             // PP_RS255_LFSR=1: the round-1 LFSR kernels instead of the transform kernels (A/B runs)
    // opt-in per-kernel timing (poporon_cuda_rs_profile): CUDA event pairs around each launch
    bool prof_on;
    int prof_used;
    cudaEvent_t *prof_ev;     // [2 * kProfCap]
    uint8_t *prof_id;         // [kProfCap]
    double prof_ms[ppcu::PROF_KINDS];
    uint64_t prof_n[ppcu::PROF_KINDS];
};

struct poporon_cuda_bch {
    int device;
    int sm_count;
    int m, nn, n, k, parity_bits, t;
    uint32_t gen;
    uint8_t *d_tables; // alpha_to[256] | index_of[256]
    uint32_t *d_map;   // n <= 15: the complete decode map, one word per received word (bch.cu), else NULL
    uint32_t *d_smap;  // n > 15, parity_bits <= 20: decode map indexed by the remainder mod g (bch.cu), else NULL
    uint64_t launches;
};
