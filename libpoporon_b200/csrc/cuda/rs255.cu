// rs255.cu — bitsliced sm_100a kernels for the reference's default code RS(255,223), GF(2^8)/0x11D,
// fcr = 1, prim = 1 (reference src/poporon.c:281-284).
//
// What replaces what (paths relative to the reference tree):
//   rs_encode (src/encode.c:17-145) ............ rs255_lfsr_kernel<ENCODE>: systematic parity
//   calculate_syndrome_u8 (src/decode.c:232-416)  rs255_lfsr_kernel<SCREEN>: "is the codeword clean" + the
//                                                 32 syndromes of dirty codewords
//   error_correction_u8 (src/decode.c:17-230) ... rs255_correct_kernel on a compacted queue of dirty
//                                                 codewords (rs_correct.cuh)
//
// Data layout.  A bit-plane is one 32-bit register holding the same bit of the same symbol of 32
// codewords.  A multiplication by a field CONSTANT is then an 8x8 GF(2) matrix = a fixed XOR network,
// and the LFSR step "state <- state * x + fb * g" for 32 codewords costs one LOP3 (3-input XOR) per state
// bit: every state bit is updated with a subset-XOR of the 8 feedback planes, looked up in two 16-entry
// tables of 4-bit subset XORs that live in registers (the four-Russians trick; the subset is a
// compile-time constant from gf256_consts.cuh).
//
// 256 state planes (32 parity symbols x 8 bits) exceed one thread's registers, so g(x) is split into two
// degree-16 factors and each warp of a CTA (a "role") runs the LFSR of one factor on the same 1024
// codewords (lane l owns codewords {tile*1024 + 32*i + l}).  See gf256_consts.cuh for the algebra.
//
// Memory.  Codeword streams are fetched with 16-byte cp.async into a 3-stage shared-memory ring
// (one 16-byte chunk per codeword per stage) shared by the two roles, one round = 16 symbols ahead of
// use; a lane's 32 codewords share the same byte alignment inside their chunks, so re-alignment is one
// PRMT per word with a loop-invariant selector.  Bytes -> bit-planes is a 32x32 bit-matrix transpose in
// registers (2 PRMT stages + 3 shift/LOP3 stages).
//
// The kernels are bound by the integer ALU pipe (LOP3/PRMT at 64 lanes/clk/SM measured, tools/ubench_pipes.cu),
// not by HBM: DESIGN.md has the op count per symbol.
#include "rs255.cuh"

#include <utility>

#include "gf256_consts.cuh"
#include "rs_correct.cuh"
#include "rs255_solve_core.h"
#include "rs255_erasure_core.h"
#include "rs255_fft.cuh"
#include <stdlib.h>

namespace ppcu {

namespace {

constexpr int kMsg = 223, kPar = 32;
constexpr int kTile = 1024;              // codewords per CTA tile

enum { MODE_ENCODE = 0, MODE_SCREEN = 1 };

template <int ROLE>
struct Masks {
    static constexpr gf256::LfsrMasks<ROLE> lfsr = gf256::make_lfsr_masks<ROLE>();
};

// encoder recombination: output half OH (parity bytes 16*OH .. 16*OH+15) from the state of role SRC
template <int SRC>
struct CrtM {
    static constexpr gf256::CrtMasks<SRC> v = gf256::make_crt_masks<SRC>();
};

// syndrome map of a role: the role's register holds s = c(x) * x^16 mod g_role (coefficient k of s = x^(15-k)), so
// for a root alpha^j of g_role, j = 1 + 16*ROLE + q:  S_j = c(alpha^j) = s(alpha^j) * alpha^(-16 j)
//                                                          = sum_k s[k] * alpha^(j*(15-k) - 16 j) = sum_k s[k] * alpha^(-j*(k+1))
template <int ROLE>
struct SynM {
    uint8_t m[16][16][8];
};
template <int ROLE>
constexpr SynM<ROLE> make_synm()
{
    SynM<ROLE> r{};
    for (int k = 0; k < 16; k++)
        for (int q = 0; q < 16; q++)
            for (int t = 0; t < 8; t++)
                r.m[k][q][t] = gf256::rowmask(gf256::pow_alpha(-(1 + 16 * ROLE + q) * (k + 1)), t);
    return r;
}
template <int ROLE>
struct SynMasks {
    static constexpr SynM<ROLE> v = make_synm<ROLE>();
};

// ---------------------------------------------------------------------------------------------------
// small device helpers
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void cp_async16(uint32_t saddr, const void *gaddr, uint32_t src_bytes)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(saddr), "l"(gaddr), "r"(src_bytes)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit()
{
    asm volatile("cp.async.commit_group;\n" ::: "memory");
}
__device__ __forceinline__ void cp_async_wait_all()
{
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
}
__device__ __forceinline__ uint32_t lds32(uint32_t saddr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];\n" : "=r"(v) : "r"(saddr));
    return v;
}
__device__ __forceinline__ void sts32(uint32_t saddr, uint32_t v)
{
    asm volatile("st.shared.u32 [%0], %1;\n" ::"r"(saddr), "r"(v) : "memory");
}

// (a & m) | (b & ~m) in ONE LOP3 (immLut 0xE4).  Written as PTX because the optimiser otherwise "unfolds" the
// masked merge into ((a ^ b) & m) ^ b plus separate ANDs: 4+ instructions per word instead of 1.
__device__ __forceinline__ uint32_t merge_bits(uint32_t a, uint32_t b, uint32_t m)
{
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xE4;" : "=r"(d) : "r"(a), "r"(b), "r"(m));
    return d;
}

// 32x32 bit-matrix transpose in registers: x[r] bit c  <->  x[c] bit r.  An involution, used in both
// directions (bytes -> planes on the way in, planes -> bytes on the way out).
__device__ __forceinline__ void transpose32(uint32_t (&x)[32])
{
#pragma unroll
    for (int r = 0; r < 16; r++) { // column bit 4 <-> row bit 4: half-word exchange
        const uint32_t a = x[r], b = x[r + 16];
        x[r] = __byte_perm(a, b, 0x5410);
        x[r + 16] = __byte_perm(a, b, 0x7632);
    }
#pragma unroll
    for (int r = 0; r < 32; r++) { // bit 3: byte exchange
        if (r & 8)
            continue;
        const uint32_t a = x[r], b = x[r + 8];
        x[r] = __byte_perm(a, b, 0x6240);
        x[r + 8] = __byte_perm(a, b, 0x7351);
    }
#pragma unroll
    for (int r = 0; r < 32; r++) { // bit 2: nibble exchange
        if (r & 4)
            continue;
        const uint32_t a = x[r], b = x[r + 4];
        x[r] = merge_bits(a, b << 4, 0x0F0F0F0Fu);
        x[r + 4] = merge_bits(a >> 4, b, 0x0F0F0F0Fu);
    }
#pragma unroll
    for (int r = 0; r < 32; r++) { // bit 1
        if (r & 2)
            continue;
        const uint32_t a = x[r], b = x[r + 2];
        x[r] = merge_bits(a, b << 2, 0x33333333u);
        x[r + 2] = merge_bits(a >> 2, b, 0x33333333u);
    }
#pragma unroll
    for (int r = 0; r < 32; r++) { // bit 0
        if (r & 1)
            continue;
        const uint32_t a = x[r], b = x[r + 1];
        x[r] = merge_bits(a, b << 1, 0x55555555u);
        x[r + 1] = merge_bits(a >> 1, b, 0x55555555u);
    }
}

// a ^ b ^ c as ONE LOP3, spelled in PTX: given plain C xors ptxas re-associates "P ^ T1[a] ^ T2[b]" into chains
// over the input planes (it does not see that the table entries are shared) and spends 213 instead of 158
// instructions per LFSR step
__device__ __forceinline__ uint32_t xor3(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// all 16 subset-XORs of four planes: T[s] = xor of b_k over the bits k set in s (11 XORs)
__device__ __forceinline__ void subset_table(uint32_t b0, uint32_t b1, uint32_t b2, uint32_t b3, uint32_t (&T)[16])
{
    T[0] = 0;
    T[1] = b0;
    T[2] = b1;
    T[3] = b0 ^ b1;
    T[4] = b2;
    T[5] = b2 ^ b0;
    T[6] = b2 ^ b1;
    T[7] = b2 ^ T[3];
    T[8] = b3;
    T[9] = b3 ^ b0;
    T[10] = b3 ^ b1;
    T[11] = b3 ^ T[3];
    T[12] = b3 ^ b2;
    T[13] = T[12] ^ b0;
    T[14] = T[12] ^ b1;
    T[15] = T[12] ^ T[3];
}

// ---- one LFSR step for 32 codewords: 8 + 22 + 128 three-input XORs --------------------------------
template <int ROLE, int JT>
__device__ __forceinline__ void lfsr_update_one(uint32_t (&P)[16][8], const uint32_t (&T1)[16],
                                                const uint32_t (&T2)[16])
{
    constexpr int J = JT >> 3, T = JT & 7;
    constexpr unsigned m = Masks<ROLE>::lfsr.m[J][T];
    if constexpr (J < 15)
        P[J][T] = xor3(P[J + 1][T], T1[m & 15], T2[m >> 4]);
    else
        P[J][T] = T1[m & 15] ^ T2[m >> 4];
}

template <int ROLE, int... JT>
__device__ __forceinline__ void lfsr_update_all(uint32_t (&P)[16][8], const uint32_t (&T1)[16],
                                                const uint32_t (&T2)[16], std::integer_sequence<int, JT...>)
{
    (lfsr_update_one<ROLE, JT>(P, T1, T2), ...);
}

template <int ROLE>
__device__ __forceinline__ void lfsr_step(uint32_t (&P)[16][8], const uint32_t *d /* 8 planes of the symbol */)
{
    uint32_t T1[16], T2[16];
    subset_table(d[0] ^ P[0][0], d[1] ^ P[0][1], d[2] ^ P[0][2], d[3] ^ P[0][3], T1);
    subset_table(d[4] ^ P[0][4], d[5] ^ P[0][5], d[6] ^ P[0][6], d[7] ^ P[0][7], T2);
    lfsr_update_all<ROLE>(P, T1, T2, std::make_integer_sequence<int, 128>{});
}

// ---- constant linear map on bit-planes: acc[q] ^= in * c[k][q] for 16 outputs ------------------------
template <class M, int K, int OBASE, int QT>
__device__ __forceinline__ void linmap_one(uint32_t (&acc)[16][8], const uint32_t (&T1)[16], const uint32_t (&T2)[16])
{
    constexpr int Q = QT >> 3, T = QT & 7;
    constexpr unsigned m = M::v.m[K][OBASE + Q][T];
    acc[Q][T] = xor3(acc[Q][T], T1[m & 15], T2[m >> 4]);
}
template <class M, int K, int OBASE, int... QT>
__device__ __forceinline__ void linmap_all(uint32_t (&acc)[16][8], const uint32_t (&T1)[16], const uint32_t (&T2)[16],
                                           std::integer_sequence<int, QT...>)
{
    (linmap_one<M, K, OBASE, QT>(acc, T1, T2), ...);
}
// input symbol k of a 16-symbol vector stored in shared memory as [plane = 8k + t][lane] words
template <class M, int OBASE, int K>
__device__ __forceinline__ void linmap_k(uint32_t (&acc)[16][8], uint32_t sbase /* smem addr of plane 0, this lane */)
{
    uint32_t in[8], T1[16], T2[16];
#pragma unroll
    for (int t = 0; t < 8; t++)
        in[t] = lds32(sbase + (uint32_t)((8 * K + t) * 128));
    subset_table(in[0], in[1], in[2], in[3], T1);
    subset_table(in[4], in[5], in[6], in[7], T2);
    linmap_all<M, K, OBASE>(acc, T1, T2, std::make_integer_sequence<int, 128>{});
}
template <class M, int OBASE, int... K>
__device__ __forceinline__ void linmap_vec(uint32_t (&acc)[16][8], uint32_t sbase, std::integer_sequence<int, K...>)
{
    (linmap_k<M, OBASE, K>(acc, sbase), ...);
}

// ---------------------------------------------------------------------------------------------------
// the stream kernel: one CTA = two warps (roles) x one tile of 1024 codewords at a time
// ---------------------------------------------------------------------------------------------------
#ifndef PP_LFSR_PAIRS
#define PP_LFSR_PAIRS 4
#endif
constexpr int kPairs = PP_LFSR_PAIRS; // role pairs (64 threads each) per CTA: 4, 2 or 1

// Which warp of the CTA plays which role of which pair.  A scheduler (SM sub-partition) owns warps w and w + 4.
// 0: pair = warp / 2, role = warp % 2          (each scheduler: the same role of two pairs)
// 1: as 0, roles swapped in pairs 2 and 3      (each scheduler: one warp of either role)
// 2: pair = warp % 4, role = warp / 4          (each scheduler: both roles of one pair)
#ifndef PP_LFSR_PREFETCH
#define PP_LFSR_PREFETCH 1
#endif
#ifndef PP_ROLE_MAP
#define PP_ROLE_MAP 1
#endif
__device__ __forceinline__ int lfsr_pair()
{
    return PP_ROLE_MAP == 2 ? (int)((threadIdx.x >> 5) & 3u) : (int)(threadIdx.x >> 6);
}
__device__ __forceinline__ int lfsr_role()
{
    const unsigned w = threadIdx.x >> 5;
    return PP_ROLE_MAP == 2 ? (int)(w >> 2) : PP_ROLE_MAP == 1 ? (int)((w ^ (w >> 2)) & 1u) : (int)(w & 1u);
}

// barrier of one role pair (64 threads); barrier 0 is the CTA-wide __syncthreads()
__device__ __forceinline__ void pair_sync(int pair)
{
    asm volatile("bar.sync %0, 64;" ::"r"(pair + 1) : "memory");
}

// Tile geometry.  Codewords are taken in blocks of 4096; a block is four tiles ("groups" G = 0..3).  Lane l of
// the tile owns the 32 codewords  cw = 4096*block + 512*(l >> 2) + 16*i + r,  i = 0..31, where r is one of the
// four residues (mod 16) of the group.  Consecutive i are 16 codewords apart, so all 32 streams of a lane
// have the same byte offset o inside their 16-byte chunks whatever the stride is; and the host sorts the 16
// residues by (o >> 2) before cutting them into groups, so (o >> 2) - the word offset - is the same for the
// WHOLE tile.  That makes the chunk schedule uniform per CTA and a two-stage ring sufficient.
struct LfsrArgs {
    const uint8_t *stream; // first byte of codeword 0's stream (data for ENCODE, packed record for SCREEN)
    size_t stride;         // bytes between consecutive codewords' streams
    int len;               // symbols fed to the LFSR (size, or size + 32 for SCREEN)
    size_t n;              // codewords
    const uint8_t *gend;   // one past the last readable byte of the stream buffer
    uint64_t perm;         // nibble p: residue (codeword index mod 16) in slot p; group G = slots 4G..4G+3
    uint32_t wbs;          // 2 bits per group: word offset (o >> 2) of the group's streams
    // ENCODE outputs
    uint8_t *parity;
    size_t parity_stride;
    uint8_t *ok;
    // SCREEN outputs
    uint32_t *dmask; // [2 * groups]: per (tile, lane, role) 32-bit "state != 0" mask over i
    uint8_t *syn;    // [n][32]: the syndromes S_1 .. S_32 (field elements), 16 per role
};

constexpr int kRingStage = 16384;          // [i = 32][lane = 32][16 B]
constexpr int kRingBytes2 = 2 * kRingStage; // two stages
constexpr int kWinBytes = 4096;            // one window of 4 symbols: [plane = 32][lane = 32] words
constexpr int kSmemEncode = kRingBytes2 + 6 * kWinBytes;
constexpr int kSmemScreen = kRingBytes2 + 4 * kWinBytes;

__device__ __forceinline__ size_t tile_cw0(const LfsrArgs &A, size_t tile, int lane)
{
    const unsigned slot = 4u * (unsigned)(tile & 3) + (unsigned)(lane & 3);
    const unsigned r = (unsigned)(A.perm >> (4u * slot)) & 15u;
    return (tile >> 2) * 4096 + 512u * (unsigned)(lane >> 2) + r;
}

// one update of a role's 16-symbol register with the multiplier symbol u (8 planes):
// P[j] <- P[j+1] ^ u * g_{15-j}
template <int ROLE>
__device__ __forceinline__ void lfsr_update(uint32_t (&P)[16][8], const uint32_t (&u)[8])
{
    uint32_t T1[16], T2[16];
    subset_table(u[0], u[1], u[2], u[3], T1);
    subset_table(u[4], u[5], u[6], u[7], T2);
    lfsr_update_all<ROLE>(P, T1, T2, std::make_integer_sequence<int, 128>{});
}

// KIND 0: divider      u = d ^ P[0]                       (screen roles, encoder role 1)
// KIND 1: divider that also emits the quotient symbol e = d ^ P[0] in place of d   (encoder role 0)
// KIND 2: multiplier   u = d, emits e = d ^ P[0] in place of d   (encoder role 0, last 16 steps)
// d holds the 8 planes of the symbol at saddr (loaded by the caller, one symbol ahead of use).
template <int ROLE, int KIND>
__device__ __forceinline__ void stream_step(uint32_t (&P)[16][8], const uint32_t (&d)[8], uint32_t saddr)
{
    uint32_t e[8];
#pragma unroll
    for (int t = 0; t < 8; t++)
        e[t] = d[t] ^ P[0][t];
    if (KIND != 0) {
#pragma unroll
        for (int t = 0; t < 8; t++)
            sts32(saddr + (uint32_t)(t * 128), e[t]);
    }
    if (KIND == 2)
        lfsr_update<ROLE>(P, d);
    else
        lfsr_update<ROLE>(P, e);
}

__device__ __forceinline__ void load_symbol(uint32_t (&d)[8], uint32_t saddr)
{
#pragma unroll
    for (int t = 0; t < 8; t++)
        d[t] = lds32(saddr + (uint32_t)(t * 128));
}

// `count` consecutive symbols starting at saddr (1 KB apart), the next symbol's planes loaded while the current
// one is being processed
template <int ROLE, int KIND>
__device__ __forceinline__ void stream_run(uint32_t (&P)[16][8], uint32_t saddr, int count)
{
    if (count <= 0)
        return;
    uint32_t d0[8], d1[8];
    load_symbol(d0, saddr);
#pragma unroll 1
    for (int y = 0; y < count; y += 2) {
        if (y + 1 < count)
            load_symbol(d1, saddr + 1024u);
        stream_step<ROLE, KIND>(P, d0, saddr);
        if (y + 1 < count) {
            if (y + 2 < count)
                load_symbol(d0, saddr + 2048u);
            stream_step<ROLE, KIND>(P, d1, saddr + 1024u);
        }
        saddr += 2048u;
    }
}

__device__ __forceinline__ uint2 lds64(uint32_t saddr)
{
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];\n" : "=r"(v.x), "=r"(v.y) : "r"(saddr));
    return v;
}

template <int MODE, int ROLE>
__device__ __forceinline__ void stream_role(const LfsrArgs &A, uint8_t *smem)
{
    // A step is two windows (8 symbols); window buffers come in sets of two, NB sets in rotation
    constexpr int NB = MODE == MODE_ENCODE ? 3 : 2;
    constexpr int kSetBytes = 2 * kWinBytes;
    const int lane = threadIdx.x & 31;
    const int pair = lfsr_pair(); // which of the CTA's role pairs (each works on its own tile)
    const uint32_t ring_s = smem_u32(smem);
    const uint32_t lane_ring = ring_s + (uint32_t)lane * 16u;
    const uint32_t lane_buf = ring_s + (uint32_t)kRingBytes2 + (uint32_t)lane * 4u;
    const size_t ntiles = 4 * ((A.n + 4095) / 4096);
    const int nW = (A.len + 3) >> 2;
    const int nS = (nW + 1) >> 1;

    // The four pairs of a CTA take the four tiles of one 4096-codeword block and meet at a CTA-wide barrier after
    // every round: all eight warps of the SM then run the same phase (main loop, epilogue) at the same time and
    // share its code in the instruction cache (the straight-line epilogue of a pair drifting on its own evicted the
    // main loops of the three others).
    const size_t vstride = (size_t)gridDim.x * kPairs;
    const size_t rounds = (ntiles + vstride - 1) / vstride;
    bool have0 = false; // chunk 0 of the tile was requested during the previous tile's epilogue
    for (size_t it = 0; it < rounds; it++, __syncthreads()) {
        const size_t tile = it * vstride + (size_t)blockIdx.x * kPairs + (size_t)pair;
        if (tile >= ntiles)
            continue;
        const size_t cw0 = tile_cw0(A, tile, lane);
        const int wb = (int)((A.wbs >> (2u * (unsigned)(tile & 3))) & 3u);
        const uint8_t *s0 = A.stream + cw0 * A.stride;
        const uint32_t o = (uint32_t)((uintptr_t)s0 & 15u); // (o >> 2) == wb for every lane of the tile
        const size_t istride = 16 * A.stride;
        const uint8_t *g0 = s0 - o + (size_t)(16 * ROLE) * istride; // chunk 0 of this role's first codeword
        const int nchunks = (4 * wb + 3 + A.len + 15) >> 4;
        const uint32_t sel = 0x3210u + 0x1111u * (o & 3u);
        // whole tile inside the batch and every chunk inside the buffer: no per-copy checks (all but the last tiles)
        const bool interior = ((tile >> 2) + 1) * 4096 <= A.n &&
                              A.stream + (((tile >> 2) + 1) * 4096 - 1) * A.stride + (size_t)nchunks * 16 + 16 <= A.gend;

        // this role fetches chunk c of its half of the codewords (i in [16*ROLE, 16*ROLE + 16)) of a tile
        auto fetch_from = [&](const uint8_t *tg0, int tnchunks, bool tinterior, size_t tcw0, int c) {
            if (c < tnchunks) {
                const uint32_t dst = lane_ring + (uint32_t)((c & 1) * kRingStage + 16 * ROLE * 512);
                const uint8_t *g = tg0 + (size_t)c * 16;
                if (tinterior) {
#pragma unroll
                    for (int ii = 0; ii < 16; ii++)
                        cp_async16(dst + (uint32_t)(ii * 512), g + (size_t)ii * istride, 16u);
                } else {
#pragma unroll 1
                    for (int ii = 0; ii < 16; ii++) {
                        const uint8_t *gi = g + (size_t)ii * istride;
                        uint32_t nbytes = 0;
                        if (tcw0 + 16 * (size_t)(16 * ROLE + ii) < A.n && gi < A.gend) {
                            const size_t left = (size_t)(A.gend - gi);
                            nbytes = left >= 16 ? 16u : (uint32_t)left;
                        }
                        cp_async16(dst + (uint32_t)(ii * 512), nbytes ? gi : A.stream, nbytes);
                    }
                }
            }
            cp_async_commit();
        };
        auto fetch = [&](int c) { fetch_from(g0, nchunks, interior, cw0, c); };
        // chunk 0 of this pair's NEXT tile, issued when the main loop is over: ring stage 0 is kept out of the
        // epilogue's scratch, so the copy (32 different lines per instruction: slow to issue) runs under the epilogue
        // instead of in front of an idle SM at the start of the next round
        auto prefetch_next = [&]() -> bool {
            const size_t nt = tile + vstride;
            if (nt >= ntiles)
                return false;
            const size_t ncw0 = tile_cw0(A, nt, lane);
            const int nwb = (int)((A.wbs >> (2u * (unsigned)(nt & 3))) & 3u);
            const uint8_t *ns0 = A.stream + ncw0 * A.stride;
            const uint32_t no = (uint32_t)((uintptr_t)ns0 & 15u);
            const bool ninterior = ((nt >> 2) + 1) * 4096 <= A.n &&
                                   A.stream + (((nt >> 2) + 1) * 4096 - 1) * A.stride +
                                           (size_t)((4 * nwb + 3 + A.len + 15) >> 4) * 16 + 16 <= A.gend;
            fetch_from(ns0 - no + (size_t)(16 * ROLE) * istride, (4 * nwb + 3 + A.len + 15) >> 4, ninterior, ncw0, 0);
            return true;
        };

        // bytes -> bit-planes of window W (symbols 4W .. 4W+3) into window buffer at dst (this lane's column).
        // The window is words jlo, jlo+1 of every stream; jlo is the same for the whole tile.
        auto transpose_window = [&](int W, uint32_t dst) {
            const uint32_t jlo = (uint32_t)(wb + W);
            uint32_t X[32];
            if ((jlo & 1u) == 0) { // one aligned 8-byte load per codeword
                const uint32_t off = ((jlo >> 2) & 1u) * (uint32_t)kRingStage + (jlo & 3u) * 4u;
#pragma unroll
                for (int i = 0; i < 32; i++) {
                    const uint2 v = lds64(lane_ring + (uint32_t)(i * 512) + off);
                    X[i] = __byte_perm(v.x, v.y, sel);
                }
            } else { // the two words straddle an 8-byte boundary (possibly two ring stages)
                const uint32_t jhi = jlo + 1;
                const uint32_t off_lo = ((jlo >> 2) & 1u) * (uint32_t)kRingStage + (jlo & 3u) * 4u;
                const uint32_t off_hi = ((jhi >> 2) & 1u) * (uint32_t)kRingStage + (jhi & 3u) * 4u;
#pragma unroll
                for (int i = 0; i < 32; i++) {
                    const uint32_t lo = lds32(lane_ring + (uint32_t)(i * 512) + off_lo);
                    const uint32_t hi = lds32(lane_ring + (uint32_t)(i * 512) + off_hi);
                    X[i] = __byte_perm(lo, hi, sel);
                }
            }
            transpose32(X); // X[8*y + t] = plane of bit t of symbol 4W + y
#pragma unroll
            for (int p = 0; p < 32; p++)
                sts32(dst + (uint32_t)(p * 128), X[p]);
        };

        uint32_t P[16][8];
#pragma unroll
        for (int j = 0; j < 16; j++)
#pragma unroll
            for (int t = 0; t < 8; t++)
                P[j][t] = 0;

        pair_sync(pair); // the previous tile's epilogue is done with the shared memory
        if (!have0)
            fetch(0);
        fetch(1);
        cp_async_wait_all();
        pair_sync(pair);
        if (ROLE < nW)
            transpose_window(ROLE, lane_buf + (uint32_t)(ROLE * kWinBytes)); // windows 0 and 1 -> set 0
        pair_sync(pair);

        // step s: role r transposes window 2s+2+r into the next set, then both run the registers over windows
        // 2s and 2s+1 (encoder role 1 runs one step behind role 0, on the quotient symbols role 0 left in place)
        const int nsteps = MODE == MODE_ENCODE ? nS + 1 : nS;
        int set_cur = 0, set_prev = NB - 1;
#pragma unroll 1
        for (int s = 0; s < nsteps; s++) {
            const int set_next = set_cur + 1 == NB ? 0 : set_cur + 1;
            const int v = 2 * s - 2 + wb;
            if ((v & 2) == 0 && v >= 0)
                fetch((v >> 2) + 2);
            if (2 * s + 2 + ROLE < nW)
                transpose_window(2 * s + 2 + ROLE, lane_buf + (uint32_t)(set_next * kSetBytes + ROLE * kWinBytes));
            if (MODE == MODE_ENCODE && ROLE == 1) {
                if (s >= 1) {
                    const int left = A.len - 8 * (s - 1);
                    stream_run<ROLE, 0>(P, lane_buf + (uint32_t)(set_prev * kSetBytes), left < 8 ? left : 8);
                }
            } else if (s < nS) {
                const int left = A.len - 8 * s;
                stream_run<ROLE, MODE == MODE_ENCODE ? 1 : 0>(P, lane_buf + (uint32_t)(set_cur * kSetBytes),
                                                               left < 8 ? left : 8);
            }
            if ((v & 2) == 0 && v >= 0)
                cp_async_wait_all();
            pair_sync(pair);
            set_prev = set_cur;
            set_cur = set_next;
        }
        cp_async_wait_all();
        pair_sync(pair); // every chunk has landed and been consumed: the ring (but stage 0) becomes epilogue scratch
        have0 = (PP_LFSR_PREFETCH && MODE == MODE_SCREEN) ? prefetch_next() : false; // measured: -2 % screen, +2 % encoder
        uint32_t acc[16][8];
        int out_half = ROLE; // which 16 bytes of the 32-byte output row this role writes
        if (MODE == MODE_ENCODE) {
            // parity = tB * gA + sA * x^16 (sA, tB: the final registers of roles 0 and 1).  Role 0 keeps going as
            // a multiplier by gA, fed with tB's 16 symbols: the 16 symbols it emits are parity[0..15], the register
            // it is left with is parity[16..31].
            const uint32_t scr = lane_buf; // the window sets are free now (16 KB needed, 24 KB there)
            if (ROLE == 1) {
#pragma unroll
                for (int j = 0; j < 16; j++)
#pragma unroll
                    for (int t = 0; t < 8; t++)
                        sts32(scr + (uint32_t)((8 * j + t) * 128), P[j][t]);
            }
            pair_sync(pair);
            if (ROLE == 0) {
                stream_run<0, 2>(P, scr, 16);
            }
            pair_sync(pair);
            if (ROLE == 0) {
#pragma unroll
                for (int j = 0; j < 16; j++)
#pragma unroll
                    for (int t = 0; t < 8; t++)
                        acc[j][t] = P[j][t];
                out_half = 1;
            } else {
#pragma unroll
                for (int j = 0; j < 16; j++)
#pragma unroll
                    for (int t = 0; t < 8; t++)
                        acc[j][t] = lds32(scr + (uint32_t)((8 * j + t) * 128));
                out_half = 0;
            }
            pair_sync(pair); // scr has been read: the row staging below overlaps it
        } else {
            uint32_t nz = 0;
#pragma unroll
            for (int j = 0; j < 16; j++)
#pragma unroll
                for (int t = 0; t < 8; t++)
                    nz |= P[j][t];
            A.dmask[2 * (tile * 32 + (size_t)lane) + ROLE] = nz;
            if (!__any_sync(0xFFFFFFFFu, nz != 0))
                continue; // every codeword of the tile is clean for this role: its 16 syndromes are zero
            // spill the state planes: [role][plane = 8j + t][lane], then evaluate them at the role's 16 roots
            const uint32_t st_s = ring_s + (uint32_t)(kRingStage + ROLE * 16384) + (uint32_t)lane * 4u; // stage 1 + windows
#pragma unroll
            for (int j = 0; j < 16; j++)
#pragma unroll
                for (int t = 0; t < 8; t++)
                    sts32(st_s + (uint32_t)((8 * j + t) * 128), P[j][t]);
#pragma unroll
            for (int q = 0; q < 16; q++)
#pragma unroll
                for (int t = 0; t < 8; t++)
                    acc[q][t] = 0;
            __syncwarp();
            linmap_vec<SynMasks<ROLE>, 0>(acc, st_s, std::make_integer_sequence<int, 16>{});
        }

        // planes -> bytes: four 32x32 transposes give, for every codeword i, 16 output bytes as 4 words
        uint32_t *flat = &acc[0][0];
        uint32_t B[4][32];
#pragma unroll
        for (int w = 0; w < 4; w++) {
#pragma unroll
            for (int c = 0; c < 32; c++)
                B[w][c] = flat[32 * w + c];
            transpose32(B[w]);
        }
        if (MODE == MODE_ENCODE) {
            // Stage the 32-byte parity rows in shared memory (the upper half of ring stage 1 and the window sets: 32 KB;
            // stage 0 is receiving the next tile), row = 32*i + lane, then let 8 threads write one row: a store
            // instruction then covers 8 whole rows (8-16 sectors) instead of 32 different ones.
            const uint32_t rows_s = ring_s + (uint32_t)(kRingBytes2 - 2 * kWinBytes);
#pragma unroll
            for (int i = 0; i < 32; i++)
                asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};\n" ::"r"(rows_s + (uint32_t)((32 * i + lane) * 32 + 16 * out_half)),
                             "r"(B[0][i]), "r"(B[1][i]), "r"(B[2][i]), "r"(B[3][i])
                             : "memory");
            pair_sync(pair);
            // pass 1, whole words: thread (t & 7) = s of a group of 8 writes aligned word s+1 of the row (s < 7;
            // s = 7 writes aligned word 0 when the row happens to be 4-byte aligned).  Rows of this thread:
            // 32*i + 8*k + (t >> 3), k = 0..3, i.e. codeword i of lane 8*k + (t >> 3).
            {
                const int t = 32 * ROLE + lane, sub = t & 7;
                uint8_t *gq[4];   // address of this thread's word in row i = 0
                uint32_t shv[4];  // funnel shift (bits)
                int cnt[4];       // rows (values of i) to write
                const uint32_t wsel = sub < 7 ? (uint32_t)(4 * sub) : 0u; // row byte offset of the low source word
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const size_t c0 = tile_cw0(A, tile, 8 * k + (t >> 3));
                    uint8_t *p = A.parity + c0 * A.parity_stride;
                    const uint32_t al = (uint32_t)((uintptr_t)p & 3u);
                    gq[k] = p - al + (sub < 7 ? 4 * (sub + 1) : 0);
                    shv[k] = sub < 7 ? 8u * al : 32u; // s = 7: funnel(x, w0, 32 & 31 = 0) = w0
                    const size_t rows = c0 < A.n ? (A.n - c0 + 15) / 16 : 0;
                    cnt[k] = (sub == 7 && al != 0) ? 0 : (int)(rows < 32 ? rows : 32);
                }
                const size_t pstep = 16 * A.parity_stride;
                const uint32_t hsel = sub < 7 ? 4u : 0u;
                uint32_t row_s = rows_s + (uint32_t)((t >> 3) * 32) + wsel;
                // two rows per k and trip: the 16 shared-memory loads are issued before the first store needs one
                // (with all warps of the SM in this loop at the same time nobody else hides an LDS -> STG latency)
#pragma unroll 1
                for (int i = 0; i < 32; i += 2) {
                    uint32_t lo[2][4], hi[2][4];
#pragma unroll
                    for (int u = 0; u < 2; u++)
#pragma unroll
                        for (int k = 0; k < 4; k++) {
                            lo[u][k] = lds32(row_s + (uint32_t)(u * 1024 + k * 256));
                            hi[u][k] = lds32(row_s + (uint32_t)(u * 1024 + k * 256) + hsel);
                        }
#pragma unroll
                    for (int u = 0; u < 2; u++)
#pragma unroll
                        for (int k = 0; k < 4; k++)
                            if (i + u < cnt[k])
                                *reinterpret_cast<uint32_t *>(gq[k] + (size_t)u * pstep) =
                                    __funnelshift_l(lo[u][k], hi[u][k], shv[k]);
#pragma unroll
                    for (int k = 0; k < 4; k++)
                        gq[k] += 2 * pstep;
                    row_s += 2048u;
                }
            }
            // pass 2, the partial words at both ends of a row that is not 4-byte aligned: every thread takes rows of
            // its own lane (i = 2*j + role)
            {
                uint8_t *p = A.parity + cw0 * A.parity_stride;
                const uint32_t al = (uint32_t)((uintptr_t)p & 3u), sh = 8u * al;
                if (al != 0) {
                    const size_t pstep = 16 * A.parity_stride;
                    uint8_t *q = p - al + (size_t)ROLE * pstep;
                    uint32_t row_s = rows_s + (uint32_t)((32 * ROLE + lane) * 32);
#pragma unroll 1
                    for (int i = ROLE; i < 32; i += 8) { // four rows per trip, loads first
                        uint32_t w0[4], w7[4];
#pragma unroll
                        for (int u = 0; u < 4; u++) {
                            w0[u] = lds32(row_s + (uint32_t)(u * 2048));
                            w7[u] = lds32(row_s + (uint32_t)(u * 2048) + 28u);
                        }
#pragma unroll
                        for (int u = 0; u < 4; u++) {
                            if (cw0 + 16 * (size_t)(i + 2 * u) < A.n) {
                                uint8_t *qq = q + (size_t)(2 * u) * pstep;
                                qq[3] = (uint8_t)(w0[u] >> (24u - sh));
                                if (al <= 2)
                                    qq[2] = (uint8_t)(w0[u] >> (16u - sh));
                                if (al <= 1)
                                    qq[1] = (uint8_t)w0[u];
                                const uint32_t tail = w7[u] >> (32u - sh);
                                qq[32] = (uint8_t)tail;
                                if (al >= 2)
                                    qq[33] = (uint8_t)(tail >> 8);
                                if (al >= 3)
                                    qq[34] = (uint8_t)(tail >> 16);
                            }
                        }
                        q += 8 * pstep;
                        row_s += 8192u;
                    }
                }
            }
        } else {
#pragma unroll
            for (int i = 0; i < 32; i++) {
                const size_t cw = cw0 + 16 * (size_t)i;
                if (cw >= A.n)
                    continue;
                *reinterpret_cast<uint4 *>(A.syn + cw * 32 + 16 * ROLE) = make_uint4(B[0][i], B[1][i], B[2][i], B[3][i]);
            }
        }
    }
}

template <int MODE>
__global__ void __launch_bounds__(64 * kPairs, 4 / kPairs) rs255_lfsr_kernel(LfsrArgs A)
{
    extern __shared__ __align__(16) uint8_t ring[];
    uint8_t *mine = ring + lfsr_pair() * (MODE == MODE_ENCODE ? kSmemEncode : kSmemScreen);
    if (lfsr_role() == 0)
        stream_role<MODE, 0>(A, mine);
    else
        stream_role<MODE, 1>(A, mine);
}


// ---------------------------------------------------------------------------------------------------
// compaction: clean codewords get their status here and cost nothing further; dirty ones are queued
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
rs255_compact_kernel(const uint32_t *__restrict__ dmask, size_t n, uint64_t invperm, uint32_t *__restrict__ queue,
                     uint32_t *__restrict__ counter, uint8_t *__restrict__ ok, uint32_t *__restrict__ corrected)
{
    const size_t cw = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t flags = 0;
    if (cw < n) {
        // inverse of tile_cw0(): codeword -> (tile, lane, i)
        const uint32_t within = (uint32_t)(cw & 4095u);
        const uint32_t slot = (uint32_t)(invperm >> (4u * (within & 15u))) & 15u;
        const uint32_t q = within >> 4, i = q & 31u, l = 4u * (q >> 5) + (slot & 3u);
        const size_t tile = (cw >> 12) * 4 + (slot >> 2);
        const uint2 m = reinterpret_cast<const uint2 *>(dmask)[tile * 32 + l];
        flags = ((m.x >> i) & 1u) | (((m.y >> i) & 1u) << 1);
        if (!flags) {
            if (ok)
                ok[cw] = 1;
            if (corrected)
                corrected[cw] = 0;
        }
    }
    const uint32_t ballot = __ballot_sync(0xFFFFFFFFu, flags != 0);
    if (ballot) {
        const int lane = threadIdx.x & 31;
        uint32_t base = 0;
        if (lane == 0)
            base = atomicAdd(counter, (uint32_t)__popc(ballot));
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (flags)
            queue[base + __popc(ballot & ((1u << lane) - 1u))] = (uint32_t)cw | (flags << 30);
    }
}

// ---------------------------------------------------------------------------------------------------
// correction of the queued codewords: one thread per dirty codeword (rs_correct.cuh)
// ---------------------------------------------------------------------------------------------------
#ifndef PP_CORRECT_SYNC
#define PP_CORRECT_SYNC 1
#endif
constexpr int kCorrectSmem = SmemWork<32>::kBytes;
#if PP_CORRECT_SYNC
using PhaseSync = CtaPhaseSync;
#else
using PhaseSync = NoPhaseSync;
#endif
__global__ void __launch_bounds__(kWorkThreads)
rs255_correct_kernel(RsCode c, const RsTables *__restrict__ gtab, int mode, uint8_t *data, size_t data_stride,
                     size_t size, uint8_t *parity, size_t parity_stride, const uint32_t *__restrict__ queue,
                     const uint32_t *__restrict__ counter, const uint8_t *__restrict__ syn_raw,
                     const uint32_t *__restrict__ pos, size_t pos_stride, const uint32_t *__restrict__ cnt,
                     uint8_t *__restrict__ ok, uint32_t *__restrict__ corrected)
{
    __shared__ SmemTables T;
    extern __shared__ __align__(16) uint8_t work[]; // [array][index][thread]: no local-memory arrays
    load_tables(T, gtab);
    SmemWork<32> wk{work + threadIdx.x};
    const auto syn = wk.arr(7);
    const uint32_t total = *counter;
    // The warps of a CTA start every codeword together (one barrier per trip): the stage is ~100 KB of code and
    // warps that drift apart evict each other's loops from the instruction cache.
    for (uint32_t q0 = blockIdx.x * blockDim.x; q0 < total; q0 += gridDim.x * blockDim.x) {
        if (PP_CORRECT_SYNC)
            __syncthreads();
        const uint32_t q = q0 + threadIdx.x;
        const bool live = q < total; // threads past the end of the queue still walk through the phase barriers
        const uint32_t entry = live ? queue[q] : 0u;
        const size_t cw = entry & 0x3FFFFFFFu;
        const uint32_t flags = live ? entry >> 30 : 0u;
        // row = S_1 .. S_32 as field elements (screen kernel) -> index form
        const uint4 *raw4 = reinterpret_cast<const uint4 *>(syn_raw + cw * 32);
#pragma unroll
        for (int half = 0; half < 2; half++) {
            uint32_t w[4] = {0, 0, 0, 0};
            if (flags & (1u << half)) {
                const uint4 v = raw4[half];
                w[0] = v.x, w[1] = v.y, w[2] = v.z, w[3] = v.w;
            }
#pragma unroll
            for (int b = 0; b < 16; b++)
                syn[16 * half + b] = T.index_of[(w[b >> 2] >> (8 * (b & 3))) & 0xFFu];
        }
        uint8_t *d = data + cw * data_stride;
        uint8_t *p = parity + cw * parity_stride;
        const int16_t pad = (int16_t)(255 - (int)c.nroots - (int)size);
        uint32_t ncorr = 0;
        bool success;
        if (mode == POPORON_CUDA_MODE_ERASURE) {
            const uint32_t ec = live ? cnt[cw] : 0u;
            success = rs_correct<32, PhaseSync>(c, T, d, size, p, wk, syn, ec <= 32u ? ec : 0u, pos + cw * pos_stride,
                                                pos_stride, true, pad, ncorr, live && ec <= 32u);
        } else {
            success = rs_correct<32, PhaseSync>(c, T, d, size, p, wk, syn, 0, nullptr, 0, false, pad, ncorr, live);
        }
        if (!live)
            continue;
        if (ok)
            ok[cw] = success ? 1 : 0;
        if (corrected)
            corrected[cw] = ncorr;
    }
}


// ---------------------------------------------------------------------------------------------------
// bitsliced solver for the queued codewords (plain mode): rs255_solve_core.h, one warp per CTA, each lane
// owns 32 queue entries (slot c of lane l = entry 1024*group + 32*c + l).  Shared memory per lane:
// 256 words of syndromes (later the per-codeword root lists) + 136 words of BM scratch (later the
// evaluator terms and root counters), word-interleaved over the lanes (bank = lane: conflict-free).
// Codewords the solver declines (register length > 16 or a vanishing leading coefficient - uncorrectable
// words only) are appended to a second queue for the literal per-codeword stage.
// ---------------------------------------------------------------------------------------------------
#ifndef PP_SOLVE_WARPS_PER_CTA
#define PP_SOLVE_WARPS_PER_CTA 8
#endif
constexpr int kSolveWarpsPerCTA = PP_SOLVE_WARPS_PER_CTA; // warps of a CTA move from tile to tile together (see the kernel)
#ifndef PP_SOLVE_W_GLOBAL
#define PP_SOLVE_W_GLOBAL 0 // experiment: the work area in the L2 scratch too (frees shared memory for 16 warps per SM)
#endif
constexpr int kSolveSmem = (PP_SOLVE_W_GLOBAL ? 0 : kSolveWarpsPerCTA * bs::kWordsW * 128) + 256; // work areas + the root table (256 B)
constexpr int kSolveScratchWords = bs::kWordsS + bs::kWordsO + bs::kWordsM + 32 + (PP_SOLVE_W_GLOBAL ? bs::kWordsW : 0);  // per lane, in global memory (L2-resident)
#ifndef PP_SOLVE_WARPS_PER_SM
#define PP_SOLVE_WARPS_PER_SM 8
#endif
constexpr int kSolveWarpsPerSM = PP_SOLVE_WARPS_PER_SM;

#ifndef PP_SOLVE_LOAD_ROWS128
#define PP_SOLVE_LOAD_ROWS128 1
#endif
// Syndrome rows (32 bytes per queued codeword: S_1 .. S_32 as field elements, written by the screen kernel) ->
// bit-planes: word W of the 32 rows of a lane is transposed into the planes of S_{4W+1} .. S_{4W+4}.  A rolled loop
// (the solver's straight-line code competes with its loop bodies for the instruction cache).
__device__ __forceinline__ void solve_load_syndromes(const uint32_t (&ent)[32], const uint8_t *__restrict__ syn_raw,
                                                     const bs::Mem<32> S)
{
#if PP_SOLVE_LOAD_ROWS128
    // Half a row (16 bytes) per load: the 32 rows of a lane are read once, as two passes of 32 quad-word loads, instead
    // of eight passes of word loads that each touched every sector again (the rows of eight warps do not stay in what
    // is left of the L1 beside 200 KB of shared memory).
#pragma unroll 1
    for (int h = 0; h < 2; h++) {
        uint4 v[32];
#pragma unroll
        for (int c = 0; c < 32; c++) {
            const uint32_t e = ent[c];
            // a half whose LFSR state was zero has no row written by the screen kernel: its 16 values are 0
            const bool live = e != 0xFFFFFFFFu && ((e >> (30 + h)) & 1u);
            v[c] = live ? __ldg(reinterpret_cast<const uint4 *>(syn_raw) + (size_t)(e & 0x3FFFFFFFu) * 2 + h)
                        : make_uint4(0u, 0u, 0u, 0u);
        }
#pragma unroll 1
        for (int q = 0; q < 4; q++) {
            uint32_t X[32];
#pragma unroll
            for (int c = 0; c < 32; c++)
                X[c] = q == 0 ? v[c].x : q == 1 ? v[c].y : q == 2 ? v[c].z : v[c].w;
            bs::transpose32(X); // X[8*y + t] = plane t of S_{4W + y + 1}, W = 4 h + q
#pragma unroll
            for (int p = 0; p < 32; p++)
                S.st(32 * (4 * h + q) + p, X[p]);
        }
    }
#else
#pragma unroll 1
    for (int W = 0; W < 8; W++) {
        uint32_t X[32];
#pragma unroll
        for (int c = 0; c < 32; c++) {
            const uint32_t e = ent[c];
            // a half whose LFSR state was zero has no row written by the screen kernel: its 16 values are 0
            const bool live = e != 0xFFFFFFFFu && ((e >> (30 + (W >> 2))) & 1u);
            X[c] = live ? __ldg(reinterpret_cast<const uint32_t *>(syn_raw) + (size_t)(e & 0x3FFFFFFFu) * 8 + W) : 0u;
        }
        bs::transpose32(X); // X[8*y + t] = plane t of S_{4W + y + 1}
#pragma unroll
        for (int p = 0; p < 32; p++)
            S.st(32 * W + p, X[p]);
    }
#endif
}

// Corrections of the PREVIOUS tile, issued in 256 small portions from inside the Chien sweep of the current one
// (bs::chien_collect calls before()/after() around every location step): portion k = list slots 2*(k & 7) and
// 2*(k & 7) + 1 of codeword c = k >> 3 of this lane.  The loads of a portion (queue entry eight steps ahead, two
// bytes of locations and two of magnitudes one step ahead) are in flight while the XOR networks of the step run,
// and the corrections leave as word-wide RED.XORs (no return value, so nothing waits for them; bytes of
// neighbouring records that share a word cannot be lost).  This replaces a separate apply kernel (2.4 ms per
// 16 Mi codewords): the scattered sector read-modify-writes now overlap the ALU-bound solver.
// Records: the solver leaves, in place of a decodable codeword's raw syndrome row, bytes 0..15 = error
// locations + 1 (0 = none) and bytes 16..31 = magnitudes; queue entries of all other codewords are blanked.
#ifndef PP_HOOK_BATCH
#define PP_HOOK_BATCH 1
#endif
struct ApplyHook {
    uint32_t g0p;          // first queue index of the previous tile (0xFFFFFFFF: nothing pending) - the only state
                           // that stays live between sweeps
    const uint32_t *queue;
    uint32_t total, lane;
    const uint32_t *qprev; // queue + g0p + lane
    uint32_t nvalid;       // codewords c < nvalid of the previous tile exist for this lane
    const uint8_t *records;
    uint8_t *data, *parity; // message rows and parity rows (one span when the records are packed)
    uint32_t stride32, pstride32;
    uint32_t size;          // message bytes per codeword
    int pad;
    uint32_t entry_cur, entry_app;
    uint32_t p2, m2; // raw 16-bit loads of a portion, zero-extended by the load itself: nothing touches them before
                     // after(), so nothing waits for them early

    __device__ __forceinline__ uint32_t entry_of(uint32_t c) const
    {
        return c < nvalid ? qprev[32u * c] : 0xFFFFFFFFu;
    }
    __device__ __forceinline__ void begin()
    {
        const uint32_t first = g0p + lane;
        nvalid = 0;
        if (g0p != 0xFFFFFFFFu && first < total) {
            const uint32_t left = (total - first + 31u) / 32u;
            nvalid = left < 32u ? left : 32u;
        }
        qprev = queue + first;
        entry_cur = entry_of(0);
        p2 = 0;
        m2 = 0;
    }
    // XORs of the portion whose loads were issued one step ago.  Symbol `loc` of the codeword is byte loc - pad of
    // its message row when that is < size, else byte loc - pad - size of its parity row (the two-pointer shape of
    // reference include/poporon.h:90-91; for packed records both expressions name the same address).
    __device__ __forceinline__ void flush()
    {
        if (m2 != 0u) {
            const size_t cw = entry_app & 0x3FFFFFFFu;
            const uintptr_t dbase = (uintptr_t)data + cw * (size_t)stride32 - (size_t)(1 + pad);
            const uintptr_t pbase = (uintptr_t)parity + cw * (size_t)pstride32 - (size_t)(1 + pad) - (size_t)size;
            const uint32_t split = size + 1u + (uint32_t)pad; // list entries >= split are parity symbols
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const uint32_t mag = (m2 >> (8 * e)) & 0xFFu;
                if (mag) {
                    const uint32_t le = (p2 >> (8 * e)) & 0xFFu; // list entry = location + 1 > pad
                    const uintptr_t ua = (le >= split ? pbase : dbase) + le;
                    asm volatile("red.relaxed.gpu.global.xor.b32 [%0], %1;" ::"l"(ua & ~(uintptr_t)3),
                                 "r"(mag << (8u * (unsigned)(ua & 3u)))
                                 : "memory");
                }
            }
        }
        m2 = 0;
    }
    // The consumer of a portion's loads sits at the top of the NEXT step (the assembler otherwise hoists the
    // first use - a compare - right behind the loads and the warp waits an L2 round trip per step).
    __device__ __forceinline__ void before(int k)
    {
        flush();
        const uint32_t sp = (uint32_t)k & 7u;
        entry_app = entry_cur;
        if (entry_cur != 0xFFFFFFFFu) {
            const uint8_t *row = records + (size_t)(entry_cur & 0x3FFFFFFFu) * 32 + 2u * sp;
            asm volatile("ld.global.u16 %0, [%1];" : "=r"(p2) : "l"(row));
            asm volatile("ld.global.u16 %0, [%1+16];" : "=r"(m2) : "l"(row));
        }
        // the next codeword's queue entry is fetched straight into entry_cur one step before its first use (a
        // separate "next" register merged by a select would be read - and waited for - every step)
        if (sp == 7u)
            entry_cur = entry_of(((uint32_t)k >> 3) + 1u);
    }
    __device__ __forceinline__ void after(int) {}
    __device__ __forceinline__ void end() { flush(); }
    // Codeword-wide form for the transform-form Chien search (33 steps per tile instead of 256): step c applies the
    // record loaded by step c - 1, loads the record of codeword c and fetches the queue entry of codeword c + 1.
    uint4 wxs, wmg;
    __device__ __forceinline__ void wbegin()
    {
        begin();
        entry_app = 0xFFFFFFFFu;
    }
    __device__ __forceinline__ void wapply()
    {
        if (entry_app != 0xFFFFFFFFu) {
            const uint32_t xw[4] = {wxs.x, wxs.y, wxs.z, wxs.w}, mw[4] = {wmg.x, wmg.y, wmg.z, wmg.w};
#if PP_HOOK_BATCH
            // All sixteen addresses and words first, in registers of their own, then the sixteen predicated RED.XORs
            // back to back.  A RED holds its operand registers until the load/store unit has taken them, and one RED
            // with 32 unrelated sectors keeps that unit busy for tens of cycles: computing the next address into the
            // same registers made every correction wait for the previous one to drain.
            const size_t cw = entry_app & 0x3FFFFFFFu;
            const uintptr_t dbase = (uintptr_t)data + cw * (size_t)stride32 - (size_t)(1 + pad);
            const uintptr_t pbase = (uintptr_t)parity + cw * (size_t)pstride32 - (size_t)(1 + pad) - (size_t)size;
            const uint32_t split = size + 1u + (uint32_t)pad; // list entries >= split are parity symbols
            uintptr_t ua[16];
            uint32_t val[16];
#pragma unroll
            for (int e = 0; e < 16; e++) {
                const uint32_t le = (xw[e >> 2] >> (8 * (e & 3))) & 0xFFu; // list entry = location + 1 > pad (0: none)
                const uint32_t mag = (mw[e >> 2] >> (8 * (e & 3))) & 0xFFu;
                const uintptr_t a = (le >= split ? pbase : dbase) + le;
                val[e] = mag << (8u * (unsigned)(a & 3u));
                ua[e] = a & ~(uintptr_t)3;
            }
#pragma unroll
            for (int e = 0; e < 16; e++)
                asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %1, 0;\n\t@p red.relaxed.gpu.global.xor.b32 [%0], %1;\n\t}"
                             ::"l"(ua[e]), "r"(val[e])
                             : "memory");
#else
#pragma unroll
            for (int h = 0; h < 8; h++) {
                p2 = (xw[h >> 1] >> (16 * (h & 1))) & 0xFFFFu;
                m2 = (mw[h >> 1] >> (16 * (h & 1))) & 0xFFFFu;
                flush();
            }
#endif
        }
    }
    __device__ __forceinline__ void wstep(int c)
    {
        wapply();
        entry_app = c < 32 ? entry_cur : 0xFFFFFFFFu;
        if (entry_app != 0xFFFFFFFFu) {
            const uint4 *row = reinterpret_cast<const uint4 *>(records + (size_t)(entry_app & 0x3FFFFFFFu) * 32);
            wxs = row[0];
            wmg = row[1];
        }
        entry_cur = entry_of((uint32_t)c + 1u);
    }
    __device__ __forceinline__ void wend()
    {
        wapply();
        entry_app = 0xFFFFFFFFu;
    }
    // everything still pending (after a warp's last tile): whole records, four codewords per round so that the
    // exposed round trips (entry -> record -> XORs) are few
    __device__ __forceinline__ void drain()
    {
        begin();
#pragma unroll 1
        for (uint32_t c0 = 0; c0 < 32u; c0 += 4u) {
            uint32_t ent[4];
            uint4 xs[4], mg[4];
#pragma unroll
            for (int u = 0; u < 4; u++)
                ent[u] = entry_of(c0 + (uint32_t)u);
#pragma unroll
            for (int u = 0; u < 4; u++) {
                xs[u] = make_uint4(0u, 0u, 0u, 0u);
                mg[u] = make_uint4(0u, 0u, 0u, 0u);
                if (ent[u] != 0xFFFFFFFFu) {
                    const uint4 *row = reinterpret_cast<const uint4 *>(records + (size_t)(ent[u] & 0x3FFFFFFFu) * 32);
                    xs[u] = row[0];
                    mg[u] = row[1];
                }
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const uint32_t xw[4] = {xs[u].x, xs[u].y, xs[u].z, xs[u].w}, mw[4] = {mg[u].x, mg[u].y, mg[u].z, mg[u].w};
                entry_app = ent[u];
#pragma unroll
                for (int h = 0; h < 8; h++) {
                    p2 = (xw[h >> 1] >> (16 * (h & 1))) & 0xFFFFu;
                    m2 = (mw[h >> 1] >> (16 * (h & 1))) & 0xFFFFu;
                    flush();
                }
            }
        }
        g0p = 0xFFFFFFFFu;
    }
};

// Erasure mode on the plain solver.  A row whose actual errors all sit on declared erasure positions decodes, in the
// reference's erasure mode, to exactly what plain decoding gives - provided the reference's erasure path is on its
// quirk-free ground (src/decode.c:34-47,211-214: positions strictly ascending, so that the i-th root belongs to the
// i-th slot; 1 <= count <= 32; positions inside the message) and the word has at most 16 actual errors: the locator
// stays the erasure locator, every declared position is a root, magnitudes are zero where the byte was right, and
// corrected_num counts the non-zero ones.  This checks those conditions for codeword c of the lane (its roots are in
// the work area's list); everything else goes to the general stage, which restates the reference literally.
__device__ __noinline__ bool erasure_row_covers_roots(const uint32_t *pp, uint32_t count, uint32_t size, int pad,
                                                      const bs::Mem<32> S, int c, int nroots)
{
    if (count == 0u || count > 32u)
        return false;
    uint32_t bm[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u}; // declared locations (position + padding), 256 bits
    int prev = -1;
    for (uint32_t j = 0; j < count; j++) {
        const uint32_t p = pp[j];
        if (p >= size || (int)p <= prev)
            return false;
        prev = (int)p;
        const uint32_t loc = p + (uint32_t)pad;
        bm[loc >> 5] |= 1u << (loc & 31u);
    }
    for (int e = 0; e < nroots; e++) {
        const uint32_t v = S.ld_byte(bs::kListBase + 4 * c + (e >> 2), e & 3); // location + 1
        const uint32_t loc = v - 1u;
        if (v == 0u || !((bm[(loc >> 5) & 7u] >> (loc & 31u)) & 1u))
            return false;
    }
    return true;
}

// One warp per CTA, eight CTAs per SM: a warp keeps only its 25 KB work area in shared memory; the syndrome planes
// and the evaluator live in a per-warp scratch in global memory (48 KB, written and read by the same lane, so it
// stays in L2).  With one warp per scheduler the ALU pipe idled ~45 % of the time behind fixed-latency waits.
// The warps of a CTA meet at a barrier after every tile: they then run the same phase (BM, sweep, Forney) at the
// same time and share its code in the 32 KB instruction cache, which eight warps drifting apart do not.
__global__ void __launch_bounds__(32 * kSolveWarpsPerCTA, kSolveWarpsPerSM / kSolveWarpsPerCTA)
rs255_solve_kernel(int nr, size_t size, uint32_t *queue, const uint32_t *__restrict__ counter, uint8_t *syn_raw,
                   uint32_t *__restrict__ scratch, uint32_t *__restrict__ queue2, uint32_t *__restrict__ counter2,
                   uint8_t *__restrict__ ok, uint32_t *__restrict__ corrected, uint8_t *data, size_t data_stride,
                   uint8_t *parity, size_t parity_stride, const uint32_t *__restrict__ era_pos, size_t era_stride,
                   const uint32_t *__restrict__ era_cnt)
{
    extern __shared__ __align__(16) uint32_t solve_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t gw = blockIdx.x * kSolveWarpsPerCTA + (uint32_t)warp, nw = gridDim.x * kSolveWarpsPerCTA;
    uint32_t *gs = scratch + (size_t)gw * (kSolveScratchWords * 32) + lane;
#if PP_SOLVE_W_GLOBAL
    const bs::Mem<32> S{gs + (bs::kWordsS + bs::kWordsO + bs::kWordsM + 32) * 32};
#else
    const bs::Mem<32> S{solve_smem + warp * (bs::kWordsW * 32) + lane}; // the work area
#endif
    const bs::Mem<32> Syn{gs};
    const bs::Mem<32> B{gs + bs::kWordsS * 32}; // the evaluator
    const bs::Mem<32> Mk{gs + (bs::kWordsS + bs::kWordsO) * 32}; // root masks of the transform-form Chien search
    // the tile's 32 queue entries of this lane (L2-resident copy: the queue itself is 64 MB of DRAM by the time the
    // outcomes are written, and three loops below would wait for it entry by entry)
    const bs::Mem<32> Qs{gs + (bs::kWordsS + bs::kWordsO + bs::kWordsM) * 32};
    uint8_t *root_tab = reinterpret_cast<uint8_t *>(solve_smem + (PP_SOLVE_W_GLOBAL ? 0 : kSolveWarpsPerCTA * bs::kWordsW * 32));
    {
        // root_tab[0] = 0 (empty list slot), root_tab[v] = alpha^v for v = 1..255: each lane fills 8 consecutive powers
        uint32_t x = 1;
        for (int i = 0; i < 8 * lane; i++) {
            x <<= 1;
            if (x & 0x100u)
                x ^= gf256::kPoly;
        }
        for (int i = 8 * lane; i < 8 * lane + 8; i++) {
            root_tab[i] = i == 0 ? (uint8_t)0 : (uint8_t)x;
            x <<= 1;
            if (x & 0x100u)
                x ^= gf256::kPoly;
        }
    }
    __syncthreads(); // (every warp wrote the same table)
    const uint32_t total = *counter;
    const int pad = 255 - nr - (int)size; // nr = roots of the code = syndromes per codeword (32, 16 or 8)
    ApplyHook hook;
    hook.records = syn_raw;
    hook.data = data;
    hook.stride32 = (uint32_t)data_stride;
    hook.parity = parity;
    hook.pstride32 = (uint32_t)parity_stride;
    hook.size = (uint32_t)size;
    hook.pad = pad;
    hook.queue = queue;
    hook.total = total;
    hook.lane = (uint32_t)lane;
    hook.g0p = 0xFFFFFFFFu;

    const uint32_t rounds = (total + nw * 1024u - 1u) / (nw * 1024u); // the same for every warp of the CTA
    for (uint32_t it = 0; it < rounds; it++) {
        const uint32_t g0 = (it * nw + gw) * 1024u;
        if (g0 < total) {
        {
            uint32_t ent[32];
#pragma unroll
            for (int c = 0; c < 32; c++) {
                const uint32_t q = g0 + 32u * (uint32_t)c + (uint32_t)lane;
                ent[c] = q < total ? queue[q] : 0xFFFFFFFFu;
            }
#pragma unroll
            for (int c = 0; c < 32; c++)
                Qs.st(c, ent[c]);
            solve_load_syndromes(ent, syn_raw, Syn);
        }
        bs::LaneOut out;
        bs::solve_lane<32, 32>(Syn, B, Mk, S, pad, out, hook, nr); // also drains the previous tile's corrections

        // Outcome of every codeword of the lane.  What the solver declines goes to the literal stage.  A decodable
        // codeword gets an error record in place of its (now dead) raw syndrome row: bytes 0..15 the error locations
        // + 1 (0 = none), bytes 16..31 the magnitudes; rs255_apply_kernel turns the records into
        // corrections.  Queue entries of all other codewords are blanked so that the apply kernel skips them.
        uint32_t okmask = 0;
        int maxroots = 0;
#pragma unroll 1
        for (int c0 = 0; c0 < 32; c0 += 4) {
            uint32_t e4[4];
#pragma unroll
            for (int u = 0; u < 4; u++)
                e4[u] = Qs.ld(c0 + u);
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int c = c0 + u;
                const uint32_t q = g0 + 32u * (uint32_t)c + (uint32_t)lane;
                if (q >= total)
                    continue;
                const uint32_t entry = e4[u];
                int nroots = 0;
                int st = bs::classify<32>(S, out, c, nroots, nr);
                if (era_cnt != nullptr) {
                    // erasure mode: only rows the plain solver decodes AND whose errors all sit on declared, ascending
                    // erasure positions are finished here; the rest (failures included: more than 16 errors can still
                    // be erasure-decodable) goes to the general stage with its raw syndromes untouched
                    const size_t cwi = entry & 0x3FFFFFFFu;
                    if (st != bs::kOk || !erasure_row_covers_roots(era_pos + cwi * era_stride, era_cnt[cwi], (uint32_t)size,
                                                                   pad, S, c, nroots))
                        st = bs::kDeclined;
                }
                if (st == bs::kOk) {
                    okmask |= 1u << c;
                    maxroots = nroots > maxroots ? nroots : maxroots;
                    uint4 *row = reinterpret_cast<uint4 *>(syn_raw + (size_t)(entry & 0x3FFFFFFFu) * 32);
                    row[0] = make_uint4(S.ld(bs::kListBase + 4 * c), S.ld(bs::kListBase + 4 * c + 1),
                                        S.ld(bs::kListBase + 4 * c + 2), S.ld(bs::kListBase + 4 * c + 3));
                    row[1] = make_uint4(0u, 0u, 0u, 0u);
                    continue;
                }
                queue[q] = 0xFFFFFFFFu;
                if (st == bs::kDeclined) {
                    queue2[atomicAdd(counter2, 1u)] = entry;
                } else {
                    const size_t cw = entry & 0x3FFFFFFFu;
                    if (ok)
                        ok[cw] = 0;
                    if (corrected)
                        corrected[cw] = 0;
                }
            }
        }

        // Forney on four list slots at a time; the magnitude words go to the error records
        uint32_t nzc[5] = {0u, 0u, 0u, 0u, 0u};
        // (no decodable codeword of the warp has more roots - a locator of higher degree, a longer evaluator - than
        // wroots: the Horner chains start at the next even number, at least 4)
        const int wroots = (int)__reduce_max_sync(0xFFFFFFFFu, (unsigned)maxroots);
        const int teff = wroots <= 4 ? 4 : ((wroots + 1) & ~1) < (nr >> 1) ? ((wroots + 1) & ~1) : (nr >> 1);
#pragma unroll 1
        for (int w = 0; w < 4; w++) {
            if (4 * w >= wroots)
                break;
            bs::forney_slots<32, 32>(Syn, B, S, w, root_tab, nzc, teff);
            uint32_t X[32];
            bs::magnitude_words<32>(S, X);
#pragma unroll
            for (int c0 = 0; c0 < 32; c0 += 8) {
                uint32_t e8[8];
#pragma unroll
                for (int u = 0; u < 8; u++)
                    e8[u] = Qs.ld(c0 + u);
#pragma unroll
                for (int u = 0; u < 8; u++)
                    if (((okmask >> (c0 + u)) & 1u) && X[c0 + u] != 0u)
                        reinterpret_cast<uint32_t *>(syn_raw + (size_t)(e8[u] & 0x3FFFFFFFu) * 32 + 16)[w] = X[c0 + u];
            }
        }
#pragma unroll 1
        for (int c0 = 0; c0 < 32; c0 += 4) {
            uint32_t e4[4];
#pragma unroll
            for (int u = 0; u < 4; u++)
                e4[u] = Qs.ld(c0 + u);
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int c = c0 + u;
                if (!((okmask >> c) & 1u))
                    continue;
                const size_t cw = e4[u] & 0x3FFFFFFFu;
                uint32_t nnz = 0;
#pragma unroll
                for (int b = 0; b < 5; b++)
                    nnz |= ((nzc[b] >> c) & 1u) << b;
                if (ok)
                    ok[cw] = 1;
                if (corrected)
                    corrected[cw] = nnz;
            }
        }
        hook.g0p = g0; // this tile's records are final: applied during the next sweep (or by the drain below)
        }
        __syncthreads();
    }
    hook.drain();
}

// ---------------------------------------------------------------------------------------------------
// Erasure mode, bitsliced (rs255_erasure_core.h): every queued row whose declared positions are regular (1 .. 32 of
// them, strictly ascending, inside the message) and whose syndromes are explained by errors ON those positions - the
// reference's Berlekamp-Massey would see only zero discrepancies - is finished here, corrections included; everything
// else goes to queue2 for the literal stage with its raw syndromes untouched.  Same shape as the solver kernel: a lane
// owns 32 queue entries, a warp's scratch (syndrome planes, Gamma, S * Gamma, slot lists) sits in the L2-resident
// scratch, the 200-word work area in shared memory.
// ---------------------------------------------------------------------------------------------------
// The dirty queue ordered by declared count (a counting sort, 34 buckets): a lane's 32 codewords then declare about the
// same number of positions, and the kernel's loops - locator build, S * Gamma, the Forney rounds - stop at the largest
// count of the warp instead of running to 32 for everybody.
constexpr int kEraBuckets = 34;   // counts 0 .. 32, 33 = anything larger
constexpr int kCntHist = 32;      // counters: [0] dirty rows, [1] rows for the literal stage, [32, 66) histogram,
constexpr int kCntCursor = 96;    //           [96, 130) bucket cursors; [128 + 32, ..) / [224, ..): the same for the second ordering
constexpr int kCntHist2 = 160, kCntCursor2 = 200;
constexpr int kCountersBytes = 1024;
__global__ void __launch_bounds__(256) rs255_era_hist_kernel(const uint32_t *__restrict__ queue, const uint32_t *__restrict__ counter,
                                                             const uint32_t *__restrict__ era_cnt, uint32_t *__restrict__ hist)
{
    __shared__ uint32_t h[kEraBuckets];
    if (threadIdx.x < kEraBuckets)
        h[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t total = *counter;
    for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < total; q += gridDim.x * blockDim.x) {
        const uint32_t f = era_cnt[queue[q] & 0x3FFFFFFFu];
        atomicAdd(&h[f < 33u ? f : 33u], 1u);
    }
    __syncthreads();
    if (threadIdx.x < kEraBuckets && h[threadIdx.x])
        atomicAdd(&hist[threadIdx.x], h[threadIdx.x]);
}
__global__ void __launch_bounds__(256) rs255_era_scatter_kernel(const uint32_t *__restrict__ queue, const uint32_t *__restrict__ counter,
                                                                const uint32_t *__restrict__ era_cnt, const uint32_t *__restrict__ hist,
                                                                uint32_t *__restrict__ cursor, uint32_t *__restrict__ out)
{
    __shared__ uint32_t start[kEraBuckets], local[kEraBuckets], base[kEraBuckets];
    if (threadIdx.x < kEraBuckets) {
        uint32_t s0 = 0;
        for (int b = 0; b < (int)threadIdx.x; b++)
            s0 += hist[b];
        start[threadIdx.x] = s0;
    }
    const uint32_t total = *counter;
    // chunks of 1024 queue entries per trip of the block: ranks inside the chunk from shared-memory counters, one
    // reservation per bucket and chunk in the global cursors
    for (uint32_t c0 = blockIdx.x * 1024u; c0 < total; c0 += gridDim.x * 1024u) {
        if (threadIdx.x < kEraBuckets)
            local[threadIdx.x] = 0;
        __syncthreads();
        uint32_t ent[4], bk[4], rk[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t q = c0 + (uint32_t)u * 256u + threadIdx.x;
            bk[u] = 0xFFFFFFFFu;
            if (q < total) {
                ent[u] = queue[q];
                const uint32_t f = era_cnt[ent[u] & 0x3FFFFFFFu];
                bk[u] = f < 33u ? f : 33u;
                rk[u] = atomicAdd(&local[bk[u]], 1u);
            }
        }
        __syncthreads();
        if (threadIdx.x < kEraBuckets)
            base[threadIdx.x] = start[threadIdx.x] + (local[threadIdx.x] ? atomicAdd(&cursor[threadIdx.x], local[threadIdx.x]) : 0u);
        __syncthreads();
#pragma unroll
        for (int u = 0; u < 4; u++)
            if (bk[u] != 0xFFFFFFFFu)
                out[base[bk[u]] + rk[u]] = ent[u];
        __syncthreads();
    }
}

// The declared positions of one row -> its slot list (eight words, a byte per slot: location + 1, 0 = slot not
// declared).  Returns whether the row is regular: 1 .. nr positions, strictly ascending, inside the message.  Every slot
// is defined by the caller (include/poporon_cuda.h): all 32 are read, as quad-words when the rows allow.
__device__ __forceinline__ bool era_row_list(const uint32_t *__restrict__ row, bool vec_rows, uint32_t f, uint32_t size, int pad,
                                             int nr, uint32_t (&w)[8])
{
    bool good = f >= 1u && f <= (uint32_t)nr;
    int prev = -1;
    uint32_t pv[32];
    if (vec_rows) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const uint4 q = __ldg(reinterpret_cast<const uint4 *>(row) + j);
            pv[4 * j] = q.x;
            pv[4 * j + 1] = q.y;
            pv[4 * j + 2] = q.z;
            pv[4 * j + 3] = q.w;
        }
    } else {
#pragma unroll
        for (int j = 0; j < 32; j++)
            pv[j] = __ldg(row + j);
    }
#pragma unroll
    for (int j = 0; j < 32; j++) {
        const uint32_t pj = pv[j];
        const bool act = (uint32_t)j < f;
        good = good && (!act || (pj < size && (int)pj > prev));
        prev = act ? (int)pj : prev;
        const uint32_t v = act ? ((pj + (uint32_t)pad + 1u) & 0xFFu) : 0u;
        w[j >> 2] |= v << (8 * (j & 3));
    }
    return good;
}
// root_tab[v] = alpha^v (the root of the locator that belongs to list byte v = location + 1), inv_tab[v] = alpha^(255 - v)
// (the X of that location); [0] = 0 in both: an undeclared slot.  Each lane fills eight consecutive powers.
__device__ __forceinline__ void era_fill_tables(uint8_t *root_tab, uint8_t *inv_tab, int lane)
{
    uint32_t x = 1;
    for (int i = 0; i < 8 * lane; i++) {
        x <<= 1;
        if (x & 0x100u)
            x ^= gf256::kPoly;
    }
    for (int i = 8 * lane; i < 8 * lane + 8; i++) { // x = alpha^i
        if (i != 0) {
            root_tab[i] = (uint8_t)x;
            if (i != 255)
                inv_tab[255 - i] = (uint8_t)x;
        }
        x <<= 1;
        if (x & 0x100u)
            x ^= gf256::kPoly;
    }
    if (lane == 0) {
        root_tab[0] = 0;
        inv_tab[0] = 0;
        inv_tab[255] = 1; // alpha^0
    }
}

// Twelve warps per SM with the work area in the scratch as well: without 200 KB of shared memory the L1 holds a good
// part of the lanes' state (8 warps: 10.3 -> 9.3 ms per 16 Mi rows), and a third warp per scheduler hides more of the
// loads that remain (9.0 ms; sixteen warps at 128 registers spill: 9.3).
#ifndef PP_ERA_WARPS
#define PP_ERA_WARPS 12
#endif
#ifndef PP_ERA_W_GLOBAL
#define PP_ERA_W_GLOBAL 1
#endif
constexpr int kEraWarps = PP_ERA_WARPS; // warps per CTA = per SM
constexpr int kEraScratchWords = bs::kWordsS + bs::kWordsG + bs::kWordsEL + 32 + (PP_ERA_W_GLOBAL ? bs::kWordsW : 0); // (S * Gamma replaces the syndromes in place)
constexpr int kEraSmem = (PP_ERA_W_GLOBAL ? 0 : kEraWarps * bs::kWordsW * 128) + 512; // work areas + the two tables
__global__ void __launch_bounds__(32 * kEraWarps, 1)
rs255_erasure_kernel(int nr, size_t size, const uint32_t *__restrict__ queue, const uint32_t *__restrict__ counter,
                     const uint8_t *__restrict__ syn_raw, uint32_t *__restrict__ scratch, uint32_t *__restrict__ queue2,
                     uint32_t *__restrict__ counter2, uint8_t *__restrict__ ok, uint32_t *__restrict__ corrected,
                     uint8_t *data, size_t data_stride, const uint32_t *__restrict__ era_pos, size_t era_stride,
                     const uint32_t *__restrict__ era_cnt)
{
    extern __shared__ __align__(16) uint32_t solve_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t gw = blockIdx.x * kEraWarps + (uint32_t)warp, nw = gridDim.x * kEraWarps;
    uint32_t *gs = scratch + (size_t)gw * (kEraScratchWords * 32) + lane;
#if PP_ERA_W_GLOBAL
    const bs::Mem<32> S{gs + (bs::kWordsS + bs::kWordsG + bs::kWordsEL + 32) * 32};
#else
    const bs::Mem<32> S{solve_smem + warp * (bs::kWordsW * 32) + lane};
#endif
    const bs::Mem<32> Syn{gs};
    const bs::Mem<32> G{gs + bs::kWordsS * 32};
    const bs::Mem<32> EL{gs + (bs::kWordsS + bs::kWordsG) * 32};
    const bs::Mem<32> Qs{gs + (bs::kWordsS + bs::kWordsG + bs::kWordsEL) * 32};
    // root_tab[v] = alpha^v (the root of Gamma that belongs to list byte v = location + 1), inv_tab[v] = alpha^(255 - v)
    // (the X of that location); [0] = 0 in both: an undeclared slot
    uint8_t *root_tab = reinterpret_cast<uint8_t *>(solve_smem + (PP_ERA_W_GLOBAL ? 0 : kEraWarps * bs::kWordsW * 32));
    uint8_t *inv_tab = root_tab + 256;
    era_fill_tables(root_tab, inv_tab, lane);
    __syncthreads(); // (every warp wrote the same tables)
    const uint32_t total = *counter;
    const int pad = 255 - nr - (int)size; // nr = roots of the code = syndromes per row = the most slots a row may declare
    const bool vec_rows = (((uintptr_t)era_pos | (era_stride * 4u)) & 15u) == 0;
    const uint32_t rounds = (total + nw * 1024u - 1u) / (nw * 1024u);
    for (uint32_t it = 0; it < rounds; it++) {
        const uint32_t g0 = (it * nw + gw) * 1024u;
        if (g0 < total) {
            {
                uint32_t ent[32];
#pragma unroll
                for (int c = 0; c < 32; c++) {
                    const uint32_t q = g0 + 32u * (uint32_t)c + (uint32_t)lane;
                    ent[c] = q < total ? queue[q] : 0xFFFFFFFFu;
                }
#pragma unroll
                for (int c = 0; c < 32; c++)
                    Qs.st(c, ent[c]);
                solve_load_syndromes(ent, syn_raw, Syn);
            }
            // the declared positions of every codeword -> slot lists (location + 1 per byte); regular rows only
            uint32_t regular = 0, present = 0;
            int maxf = 0;
#pragma unroll 1
            for (int c = 0; c < 32; c++) {
                const uint32_t entry = Qs.ld(c);
                uint32_t w[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
                bool good = false;
                if (entry != 0xFFFFFFFFu) {
                    present |= 1u << c;
                    const size_t cw = entry & 0x3FFFFFFFu;
                    const uint32_t f = era_cnt[cw];
                    const uint32_t *row = era_pos + cw * era_stride;
                    good = era_row_list(row, vec_rows, f, (uint32_t)size, pad, nr, w);
                    if (good)
                        maxf = (int)f > maxf ? (int)f : maxf;
                }
#pragma unroll
                for (int k = 0; k < 8; k++)
                    EL.st(8 * c + k, good ? w[k] : 0u);
                regular |= good ? 1u << c : 0u;
            }
            // (the queue is ordered by declared count: the warp's largest count bounds every loop below)
            const int wmaxf = (int)__reduce_max_sync(0xFFFFFFFFu, (unsigned)maxf);
            const uint32_t bad = wmaxf ? bs::erasure_lane<32, 32>(Syn, EL, G, S, inv_tab, wmaxf, nr) : 0xFFFFFFFFu;
            const uint32_t accept = regular & ~bad;
            // what does not qualify goes to the literal stage (one reservation per lane)
            {
                uint32_t rest = present & ~accept;
                if (rest) {
                    uint32_t at = atomicAdd(counter2, (uint32_t)__popc(rest));
                    while (rest) {
                        const int c = __ffs((int)rest) - 1;
                        rest &= rest - 1;
                        queue2[at++] = Qs.ld(c);
                    }
                }
            }
            uint32_t nz[6] = {0u, 0u, 0u, 0u, 0u, 0u};
            // (no row of the warp declares more than wmaxf positions: Gamma and omega end there, and so do Forney's
            // Horner chains - at the next even number, at least 4)
            const int teff = wmaxf <= 4 ? 4 : ((wmaxf + 1) & ~1) < nr ? ((wmaxf + 1) & ~1) : nr;
#pragma unroll 1
            for (int w = 0; 4 * w < nr; w++) {
                if (!__any_sync(0xFFFFFFFFu, accept != 0u && maxf > 4 * w))
                    break;
                bs::erasure_forney<32, 32>(EL, G, Syn, S, w, root_tab, nz, teff);
                uint32_t X[32];
                bs::magnitude_words<32>(S, X);
                // eight codewords per trip: their queue entries and list words (L2) are all on their way before the
                // first correction leaves
#pragma unroll 1
                for (int c0 = 0; c0 < 32; c0 += 8) {
                    uint32_t e8[8], l8[8];
#pragma unroll
                    for (int u = 0; u < 8; u++) {
                        e8[u] = Qs.ld(c0 + u);
                        l8[u] = EL.ld(8 * (c0 + u) + w);
                    }
#pragma unroll
                    for (int u = 0; u < 8; u++) {
                        const uint32_t xc = c0 == 0 ? X[u] : c0 == 8 ? X[8 + u] : c0 == 16 ? X[16 + u] : X[24 + u];
                        if (!((accept >> (c0 + u)) & 1u) || xc == 0u)
                            continue;
                        const uintptr_t base = (uintptr_t)data + (size_t)(e8[u] & 0x3FFFFFFFu) * data_stride - (size_t)(1 + pad);
#pragma unroll
                        for (int y = 0; y < 4; y++) {
                            const uintptr_t a = base + ((l8[u] >> (8 * y)) & 0xFFu);
                            const uint32_t val = ((xc >> (8 * y)) & 0xFFu) << (8u * (unsigned)(a & 3u));
                            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %1, 0;\n\t@p red.relaxed.gpu.global.xor.b32 [%0], %1;\n\t}"
                                         ::"l"(a & ~(uintptr_t)3), "r"(val)
                                         : "memory");
                        }
                    }
                }
            }
#pragma unroll 1
            for (int c = 0; c < 32; c++) {
                if (!((accept >> c) & 1u))
                    continue;
                const size_t cw = Qs.ld(c) & 0x3FFFFFFFu;
                uint32_t k = 0;
#pragma unroll
                for (int b = 0; b < 6; b++)
                    k |= ((nz[b] >> c) & 1u) << b;
                if (ok)
                    ok[cw] = 1;
                if (corrected)
                    corrected[cw] = k;
            }
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------------
// Erasure mode, second bitsliced stage (rs255_erasure_core.h, ee_lane): the rows the erasure-only kernel declined because
// their syndromes need errors OUTSIDE the declared positions.  The queue comes ordered by declared count again, and a
// warp works only when all its regular rows declare the same number F of positions (everything else - and every row
// whose sigma the plain solver would decline - goes on to the literal stage): sigma by the plain Berlekamp-Massey on
// the Forney syndromes, its roots by the plain Chien transform, then the magnitudes of the merged, location-ordered
// list of declared positions and found roots; the i-th of them is XORed into the caller's i-th SLOT, as the reference
// does (src/decode.c:211-214; slots past the row and positions past the message are skipped, as in rs_correct.cuh).
// ---------------------------------------------------------------------------------------------------
#ifndef PP_EE_B_SMEM
#define PP_EE_B_SMEM 1 // BM's B in shared memory (16 KB per warp), the rest of the work area in the scratch
#endif
constexpr int kEeWarps = 8;
constexpr int kEeScratchWords = bs::kWordsS + bs::kWordsG + bs::kWordsEL + 8 * bs::kNL + bs::kWordsM + 32 + bs::kWordsW;
static_assert(kEeWarps * kEeScratchWords <= kEraWarps * kEraScratchWords, "the erasure kernel's scratch serves both");
__global__ void __launch_bounds__(32 * kEeWarps, 1)
rs255_ee_kernel(size_t size, const uint32_t *__restrict__ queue, const uint32_t *__restrict__ counter,
                const uint8_t *__restrict__ syn_raw, uint32_t *__restrict__ scratch, uint32_t *__restrict__ queue3,
                uint32_t *__restrict__ counter3, uint8_t *__restrict__ ok, uint32_t *__restrict__ corrected, uint8_t *data,
                size_t data_stride, const uint32_t *__restrict__ era_pos, size_t era_stride,
                const uint32_t *__restrict__ era_cnt, uint32_t min_rows)
{
    __shared__ uint8_t tabs[512];
#if PP_EE_B_SMEM
    extern __shared__ __align__(16) uint32_t solve_smem[];
#endif
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t gw = blockIdx.x * kEeWarps + (uint32_t)warp, nw = gridDim.x * kEeWarps;
    uint32_t *gs = scratch + (size_t)gw * (kEeScratchWords * 32) + lane;
    const bs::Mem<32> SP{gs}; // (the block behind it, Gamma_0, is read - and not used - by the shifted BM)
    const bs::Mem<32> G{gs + bs::kWordsS * 32};
    const bs::Mem<32> EL{gs + (bs::kWordsS + bs::kWordsG) * 32};
    const bs::Mem<32> SG{gs + (bs::kWordsS + bs::kWordsG + bs::kWordsEL) * 32};
    const bs::Mem<32> MK{gs + (bs::kWordsS + bs::kWordsG + bs::kWordsEL + 8 * bs::kNL) * 32};
    const bs::Mem<32> Qs{gs + (bs::kWordsS + bs::kWordsG + bs::kWordsEL + 8 * bs::kNL + bs::kWordsM) * 32};
    const bs::Mem<32> S{gs + (bs::kWordsS + bs::kWordsG + bs::kWordsEL + 8 * bs::kNL + bs::kWordsM + 32) * 32};
    uint8_t *root_tab = tabs, *inv_tab = tabs + 256;
    era_fill_tables(root_tab, inv_tab, lane);
    __syncthreads();
    const uint32_t total = *counter;
    if (total < min_rows) {
        // A tile costs a warp the better part of a millisecond whatever the batch: with few rows the per-thread literal
        // stage is faster (64 Ki rows: 0.26 against 0.8 ms).  Everything is handed on.
        if (blockIdx.x == 0) {
            for (uint32_t q = threadIdx.x; q < total; q += blockDim.x)
                queue3[q] = queue[q];
            if (threadIdx.x == 0)
                *counter3 = total;
        }
        return;
    }
    const int pad = 255 - 32 - (int)size;
    const bool vec_rows = (((uintptr_t)era_pos | (era_stride * 4u)) & 15u) == 0;
    const uint32_t rounds = (total + nw * 1024u - 1u) / (nw * 1024u);
    for (uint32_t it = 0; it < rounds; it++) {
        const uint32_t g0 = (it * nw + gw) * 1024u;
        if (g0 < total) {
            {
                uint32_t ent[32];
#pragma unroll
                for (int c = 0; c < 32; c++) {
                    const uint32_t q = g0 + 32u * (uint32_t)c + (uint32_t)lane;
                    ent[c] = q < total ? queue[q] : 0xFFFFFFFFu;
                }
#pragma unroll
                for (int c = 0; c < 32; c++)
                    Qs.st(c, ent[c]);
                solve_load_syndromes(ent, syn_raw, SP);
            }
            uint32_t regular = 0, present = 0;
            int fmin = 64, fmax = 0;
#pragma unroll 1
            for (int c = 0; c < 32; c++) {
                const uint32_t entry = Qs.ld(c);
                uint32_t w[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
                bool good = false;
                if (entry != 0xFFFFFFFFu) {
                    present |= 1u << c;
                    const size_t cw = entry & 0x3FFFFFFFu;
                    const uint32_t f = era_cnt[cw];
                    good = era_row_list(era_pos + cw * era_stride, vec_rows, f, (uint32_t)size, pad, 32, w) && f <= 30u;
                    if (good) {
                        fmin = (int)f < fmin ? (int)f : fmin;
                        fmax = (int)f > fmax ? (int)f : fmax;
                    }
                }
#pragma unroll
                for (int k = 0; k < 8; k++)
                    EL.st(8 * c + k, good ? w[k] : 0u);
                regular |= good ? 1u << c : 0u;
            }
            const int wfmin = (int)__reduce_min_sync(0xFFFFFFFFu, (unsigned)fmin);
            const int wfmax = (int)__reduce_max_sync(0xFFFFFFFFu, (unsigned)fmax);
            uint32_t accept = 0;
            int lmax = 0;
            if (wfmax != 0 && wfmin == wfmax) { // (warp-uniform)
                const int F = wfmax;
                bs::LaneOut out;
                bs::NoHook idle;
#if PP_EE_B_SMEM
                bs::ee_lane<32, 32>(SP, EL, G, SG, MK, S, inv_tab, F, pad, out, idle, solve_smem + warp * (128 * 32) + lane);
#else
                bs::ee_lane<32, 32>(SP, EL, G, SG, MK, S, inv_tab, F, pad, out, idle);
#endif
                // outcome of sigma per codeword; the merged rank list replaces the slot list
#pragma unroll 1
                for (int c = 0; c < 32; c++) {
                    if (!((regular >> c) & 1u))
                        continue;
                    int nroots = 0;
                    int st = bs::classify<32>(S, out, c, nroots, 32 - F);
                    uint32_t m[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
                    if (st == bs::kOk) {
                        // merge (both ascending); a found root ON a declared position is a double root of the locator:
                        // the reference's search then finds fewer roots than the degree and fails
                        uint32_t lw[8], rw[4];
#pragma unroll
                        for (int k = 0; k < 8; k++)
                            lw[k] = EL.ld(8 * c + k);
#pragma unroll
                        for (int k = 0; k < 4; k++)
                            rw[k] = S.ld(bs::kListBase + 4 * c + k);
                        int a = 0, b = 0, k = 0;
                        while (a < F || b < nroots) {
                            uint32_t va = 256u, vb = 256u;
#pragma unroll
                            for (int q = 0; q < 8; q++)
                                if (q == (a >> 2) && a < F)
                                    va = (lw[q] >> (8 * (a & 3))) & 0xFFu;
#pragma unroll
                            for (int q = 0; q < 4; q++)
                                if (q == (b >> 2) && b < nroots)
                                    vb = (rw[q] >> (8 * (b & 3))) & 0xFFu;
                            if (va == vb) {
                                st = bs::kFail;
                                break;
                            }
                            const uint32_t v = va < vb ? va : vb;
                            a += va < vb ? 1 : 0;
                            b += va < vb ? 0 : 1;
#pragma unroll
                            for (int q = 0; q < 8; q++)
                                if (q == (k >> 2))
                                    m[q] |= v << (8 * (k & 3));
                            k++;
                        }
                    }
                    if (st == bs::kOk) {
                        accept |= 1u << c;
                        lmax = nroots > lmax ? nroots : lmax;
#pragma unroll
                        for (int k = 0; k < 8; k++)
                            EL.st(8 * c + k, m[k]);
                    } else {
#pragma unroll
                        for (int k = 0; k < 8; k++)
                            EL.st(8 * c + k, 0u);
                        if (st == bs::kFail) { // the reference returns false before it counts anything
                            const size_t cw = Qs.ld(c) & 0x3FFFFFFFu;
                            if (ok)
                                ok[cw] = 0;
                            if (corrected)
                                corrected[cw] = 0;
                            present &= ~(1u << c);
                        }
                    }
                }
                const int wlmax = (int)__reduce_max_sync(0xFFFFFFFFu, (unsigned)lmax);
                if (wlmax) {
                    bs::ee_products<32>(SP, G, SG, wlmax);
                    uint32_t nz[6] = {0u, 0u, 0u, 0u, 0u, 0u};
                    const int dmax = F + wlmax; // the largest degree of Gamma * sigma in the warp (<= 31)
                    const int teff = dmax <= 4 ? 4 : (dmax + 1) & ~1;
#pragma unroll 1
                    for (int w = 0; 4 * w < F + wlmax; w++) {
                        bs::erasure_forney<32, 32>(EL, G, SP, S, w, root_tab, nz, teff);
                        uint32_t X[32];
                        bs::magnitude_words<32>(S, X);
#pragma unroll 1
                        for (int c0 = 0; c0 < 32; c0 += 8) {
#pragma unroll
                            for (int u = 0; u < 8; u++) {
                                const uint32_t xc = c0 == 0 ? X[u] : c0 == 8 ? X[8 + u] : c0 == 16 ? X[16 + u] : X[24 + u];
                                if (!((accept >> (c0 + u)) & 1u) || xc == 0u)
                                    continue;
                                const size_t cw = Qs.ld(c0 + u) & 0x3FFFFFFFu;
                                const uint32_t *row = era_pos + cw * era_stride;
#pragma unroll
                                for (int y = 0; y < 4; y++) {
                                    const uint32_t mag = (xc >> (8 * y)) & 0xFFu;
                                    const size_t slot = (size_t)(4 * w + y);
                                    if (mag && slot < era_stride) { // rank -> the caller's slot of that number
                                        const uint32_t pj = __ldg(row + slot);
                                        if (pj < (uint32_t)size) {
                                            const uintptr_t a = (uintptr_t)data + cw * data_stride + pj;
                                            asm volatile("red.relaxed.gpu.global.xor.b32 [%0], %1;" ::"l"(a & ~(uintptr_t)3),
                                                         "r"(mag << (8u * (unsigned)(a & 3u)))
                                                         : "memory");
                                        }
                                    }
                                }
                            }
                        }
                    }
#pragma unroll 1
                    for (int c = 0; c < 32; c++) {
                        if (!((accept >> c) & 1u))
                            continue;
                        const size_t cw = Qs.ld(c) & 0x3FFFFFFFu;
                        uint32_t k = 0;
#pragma unroll
                        for (int b = 0; b < 6; b++)
                            k |= ((nz[b] >> c) & 1u) << b;
                        if (ok)
                            ok[cw] = 1;
                        if (corrected)
                            corrected[cw] = k;
                    }
                }
            }
            // everything else goes to the literal stage (one reservation per lane)
            {
                uint32_t rest = present & ~accept;
                if (rest) {
                    uint32_t at = atomicAdd(counter3, (uint32_t)__popc(rest));
                    while (rest) {
                        const int c = __ffs((int)rest) - 1;
                        rest &= rest - 1;
                        queue3[at++] = Qs.ld(c);
                    }
                }
            }
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------------
// Caller-supplied syndromes (reference src/decode.c:446-464: index form, nn = zero) -> the raw syndrome rows and the
// dirty queue the solver and the general stage work from.  A word is dirty when any entry differs from nn, as in
// the reference; entries above nn cannot index the reference's tables and are taken as zero (rs_generic.cu does
// the same).  Clean words are finished here (ok = 1, corrected = 0).
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
rs255_extsyn_ingest_kernel(int nr, const uint16_t *__restrict__ extsyn, size_t n, const RsTables *__restrict__ gtab,
                           uint8_t *__restrict__ syn_raw, uint32_t *__restrict__ queue, uint32_t *__restrict__ counter,
                           uint8_t *__restrict__ ok, uint32_t *__restrict__ corrected)
{
    __shared__ uint8_t exp_tab[256];
    exp_tab[threadIdx.x] = threadIdx.x < 255 ? gtab->alpha_to[threadIdx.x] : (uint8_t)0;
    __syncthreads();
    const size_t cw = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t flag = 0;
    if (cw < n) {
        const uint4 *row = reinterpret_cast<const uint4 *>(extsyn + cw * (size_t)nr); // 2 nr bytes, 16-byte aligned
        uint32_t out[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
#pragma unroll
        for (int q = 0; q < 4; q++) {
            if (8 * q >= nr)
                break; // (a code with fewer roots has shorter rows; the raw row is zero beyond them)
            const uint4 v = row[q];
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
            uint32_t b[8];
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const uint32_t sidx = (w[k >> 1] >> (16 * (k & 1))) & 0xFFFFu;
                flag |= sidx != 255u;
                b[k] = exp_tab[sidx > 255u ? 255u : sidx];
            }
            out[2 * q] = b[0] | (b[1] << 8) | (b[2] << 16) | (b[3] << 24);
            out[2 * q + 1] = b[4] | (b[5] << 8) | (b[6] << 16) | (b[7] << 24);
        }
        if (flag) {
            uint4 *dst = reinterpret_cast<uint4 *>(syn_raw + cw * 32);
            dst[0] = make_uint4(out[0], out[1], out[2], out[3]);
            dst[1] = make_uint4(out[4], out[5], out[6], out[7]);
        } else {
            if (ok)
                ok[cw] = 1;
            if (corrected)
                corrected[cw] = 0;
        }
    }
    const uint32_t ballot = __ballot_sync(0xFFFFFFFFu, flag != 0);
    if (ballot) {
        const int lane = threadIdx.x & 31;
        uint32_t base = 0;
        if (lane == 0)
            base = atomicAdd(counter, (uint32_t)__popc(ballot));
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (flag)
            queue[base + __popc(ballot & ((1u << lane) - 1u))] = (uint32_t)cw | (3u << 30);
    }
}

// The 32 syndromes of every codeword in the reference's index form (calculate_syndrome_u8, src/decode.c:232-416,
// leaves index_of[S_j], nn = 255 for zero): from the screen kernel's dirty bitmap and raw rows.  One thread per
// group of four syndromes.
__global__ void __launch_bounds__(256)
rs255_syn_export_kernel(int nr, const uint32_t *__restrict__ dirty, const uint8_t *__restrict__ syn_raw, size_t n,
                        const RsTables *__restrict__ gtab, uint16_t *__restrict__ out)
{
    __shared__ uint8_t log_tab[256];
    log_tab[threadIdx.x] = gtab->index_of[threadIdx.x];
    __syncthreads();
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t wpr = (uint32_t)nr >> 2; // groups of four syndromes per codeword
    const size_t cw = t / wpr;
    const uint32_t g = (uint32_t)(t - cw * wpr);
    if (cw >= n)
        return;
    uint32_t v = 0;
    if ((dirty[cw >> 5] >> (cw & 31u)) & 1u)
        v = reinterpret_cast<const uint32_t *>(syn_raw + cw * 32)[g];
    const uint32_t a = log_tab[v & 0xFFu], b = log_tab[(v >> 8) & 0xFFu], c = log_tab[(v >> 16) & 0xFFu], d = log_tab[v >> 24];
    reinterpret_cast<uint2 *>(out + cw * (size_t)nr)[g] = make_uint2(a | (b << 16), c | (d << 16));
}

// words_per_sm: the per-lane scratch words of all the warps an SM runs in the mode of the call (the erasure kernel keeps
// more state, in more warps, than the solver: only handles that decode in erasure mode pay for it)
int ensure_scratch(poporon_cuda_rs *ctx, size_t n, size_t words_per_sm)
{
    if (n <= ctx->queue_cap && words_per_sm <= ctx->scratch_words_per_sm)
        return 0;
    if (n < ctx->queue_cap)
        n = ctx->queue_cap;
    if (words_per_sm < ctx->scratch_words_per_sm)
        words_per_sm = ctx->scratch_words_per_sm;
    cudaFree(ctx->d_queue);
    cudaFree(ctx->d_rem);
    cudaFree(ctx->d_counters);
    ctx->d_queue = nullptr;
    ctx->d_rem = nullptr;
    ctx->d_counters = nullptr;
    ctx->queue_cap = 0;
    ctx->scratch_words_per_sm = 0;
    const size_t tiles = 4 * ((n + 4095) / 4096);
    // queue [n rounded up to 4] | dmask [tiles * 64] | queue2 [n]
    PP_CUDA_TRY(cudaMalloc((void **)&ctx->d_queue, (((n + 3) & ~(size_t)3) + tiles * 64 + n) * sizeof(uint32_t)));
    // raw syndrome rows / error records [tiles * 1024][32 B] | solver scratch, one slab per resident solver warp
    PP_CUDA_TRY(cudaMalloc((void **)&ctx->d_rem, tiles * kTile * 32 + (size_t)ctx->sm_count * words_per_sm * 128));
    PP_CUDA_TRY(cudaMalloc((void **)&ctx->d_counters, kCountersBytes));
    ctx->queue_cap = n;
    ctx->scratch_words_per_sm = words_per_sm;
    return 0;
}

} // namespace

bool rs255_matches(const poporon_cuda_rs_desc_t *d)
{
    // RS(255,223) and the reference's other two documented codes RS(255,239), RS(255,247) (README.md:417-419)
    return d->symbol_size == 8 && d->gfpoly == 0x11D && d->fcr == 1 && d->prim == 1 &&
           (d->nroots == 32 || d->nroots == 16 || d->nroots == 8);
}

// round-1 LFSR kernels: any message size 1..223; a lane's 32 codewords must be equally aligned inside their 16-byte
// chunks, which 32 * stride = 0 (mod 16) guarantees for every stride once the base is 16-byte aligned.
static bool lfsr_shape_ok(int op, const uint8_t *data, size_t data_stride, size_t size, const uint8_t *parity,
                          size_t parity_stride)
{
    if (size < 1 || size > (size_t)kMsg || data_stride < size || ((uintptr_t)data & 15u) != 0)
        return false;
    if (op == 1) // the screen kernel reads ONE contiguous stream per codeword: packed records only
        return parity == data + size && parity_stride == data_stride && data_stride <= 0xFFFFFFFFull; // 32-bit stride math
    return true;
}

// Which batches the bitsliced kernels take.  The transform kernels (rs255_fft.cu) accept any base alignment and
// the two-pointer shape of the reference API (message rows and parity rows as separate arrays, reference
// include/poporon.h:90-91) as long as 32 consecutive rows fit a sub-tile buffer; longer strides fall back to the
// LFSR kernels where those apply (16-byte aligned base, packed records for decode).
bool rs255_shape_ok(const poporon_cuda_rs *ctx, int op, const uint8_t *data, size_t data_stride, size_t size,
                    const uint8_t *parity, size_t parity_stride)
{
    const int nr = ctx->code.nroots;
    if ((!ctx->use_lfsr || nr != 32) && data_stride <= 0xFFFFFFFFull && parity_stride <= 0xFFFFFFFFull &&
        rs255_fft_shape_ok(nr, op, data, data_stride, size, parity, parity_stride, nullptr))
        return true;
    return nr == 32 && lfsr_shape_ok(op, data, data_stride, size, parity, parity_stride); // LFSR kernels: RS(255,223) only
}

int rs255_init(poporon_cuda_rs *ctx)
{
    PP_CUDA_TRY(cudaFuncSetAttribute(rs255_lfsr_kernel<MODE_ENCODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     kPairs * kSmemEncode));
    PP_CUDA_TRY(cudaFuncSetAttribute(rs255_lfsr_kernel<MODE_SCREEN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     kPairs * kSmemScreen));
    PP_CUDA_TRY(cudaFuncSetAttribute(rs255_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSolveSmem));
    PP_CUDA_TRY(cudaFuncSetAttribute(rs255_erasure_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kEraSmem));
    if (PP_EE_B_SMEM)
        PP_CUDA_TRY(cudaFuncSetAttribute(rs255_ee_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kEeWarps * 128 * 128));
    {
        const char *e = getenv("PP_RS255_NO_ERASURE_KERNEL"); // A/B: erasure mode on the plain solver + the literal stage only
        ctx->no_erasure_kernel = e && e[0] == '1';
        e = getenv("PP_RS255_NO_EE_KERNEL"); // A/B: without the second bitsliced stage of erasure mode
        ctx->no_ee_kernel = e && e[0] == '1';
        e = getenv("PP_RS255_EE_MIN_ROWS"); // rows below which that stage hands everything to the literal one (tests: 0)
        ctx->ee_min_rows = e ? (uint32_t)strtoul(e, nullptr, 10) : 131072u;
    }
    PP_CUDA_TRY(cudaFuncSetAttribute(rs255_correct_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kCorrectSmem));
    {
        int v = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, rs255_correct_kernel, kWorkThreads, kCorrectSmem) != cudaSuccess || v < 1)
            v = 4;
        ctx->correct_ctas_per_sm = v;
    }
    {
        const char *e = getenv("PP_RS255_NOFUSE"); // A/B: zero-syndrome exit as a kernel of its own
        ctx->no_fuse = e && e[0] == '1';
    }
    const char *env = getenv("PP_RS255_LFSR");
    ctx->use_lfsr = env && env[0] == '1';
    return rs255_fft_init(ctx);
}

constexpr size_t kEncChunk = (size_t)16 << 20; // codewords per encoder pass (2 x 512 MB of syndrome / parity planes)

static int ensure_slab(poporon_cuda_rs *ctx)
{
    if (ctx->d_slab)
        return 0;
    PP_CUDA_TRY(cudaMalloc((void **)&ctx->d_slab, rs255_fft_slab_bytes(ctx->sm_count)));
    return 0;
}

static unsigned lfsr_grid(const poporon_cuda_rs *ctx, size_t n)
{
    const size_t ctas = (4 * ((n + 4095) / 4096) + kPairs - 1) / kPairs; // kPairs tiles per CTA and round
    const size_t cap = (size_t)ctx->sm_count * (4 / kPairs);             // 8 warps per SM, persistent over tiles
    return (unsigned)(ctas < cap ? ctas : cap);
}

// Tile geometry for a stream buffer (see LfsrArgs): sort the 16 residues r = codeword index mod 16 by the word
// offset ((base + r*stride) mod 16) >> 2 of their streams and cut the sorted list into four groups.  Every
// offset class has 4, 8 or 16 members whatever base and stride are, so a group never straddles two classes.
static void tile_geometry(const uint8_t *stream, size_t stride, uint64_t *perm, uint64_t *invperm, uint32_t *wbs)
{
    unsigned slot = 0;
    *perm = 0, *invperm = 0, *wbs = 0;
    for (unsigned wb = 0; wb < 4; wb++)
        for (unsigned r = 0; r < 16; r++)
            if (((((uintptr_t)stream + r * stride) & 15u) >> 2) == wb) {
                *perm |= (uint64_t)r << (4 * slot);
                *invperm |= (uint64_t)slot << (4 * r);
                if ((slot & 3u) == 0)
                    *wbs |= wb << (2 * (slot >> 2));
                slot++;
            }
}

int rs255_encode(poporon_cuda_rs *ctx, const uint8_t *data, size_t data_stride, size_t size, uint8_t *parity,
                 size_t parity_stride, size_t n, uint8_t *ok, cudaStream_t stream)
{
    if (n == 0)
        return 0;
    FftShape shape;
    if ((!ctx->use_lfsr || ctx->code.nroots != 32) &&
        rs255_fft_shape_ok(ctx->code.nroots, 0, data, data_stride, size, parity, parity_stride, &shape)) {
        int rc = ensure_slab(ctx);
        if (rc != 0)
            return rc;
        // the syndrome planes between the two kernels: 32 bytes per codeword, at most kEncChunk codewords at a time
        const size_t chunk = n < kEncChunk ? n : kEncChunk;
        if (chunk > ctx->synplane_cap) {
            cudaFree(ctx->d_synplanes);
            ctx->d_synplanes = nullptr;
            ctx->synplane_cap = 0;
            PP_CUDA_TRY(cudaMalloc((void **)&ctx->d_synplanes, rs255_fft_synplane_bytes(chunk)));
            ctx->synplane_cap = chunk;
        }
        if (ok)
            PP_CUDA_TRY(cudaMemsetAsync(ok, 1, n, stream));
        for (size_t c0 = 0; c0 < n && rc == 0; c0 += chunk) {
            const size_t m = n - c0 < chunk ? n - c0 : chunk;
            rc = rs255_fft_encode(ctx, shape, data + c0 * data_stride, data_stride, size, parity + c0 * parity_stride,
                                  parity_stride, m, ctx->d_slab, ctx->d_synplanes, stream);
            ctx->launches += 3;
        }
        return rc;
    }
    LfsrArgs A{};
    A.stream = data;
    A.stride = data_stride;
    A.len = (int)size;
    A.n = n;
    A.gend = data + (n - 1) * data_stride + size;
    A.parity = parity;
    A.parity_stride = parity_stride;
    A.ok = ok;
    uint64_t invperm;
    tile_geometry(data, data_stride, &A.perm, &invperm, &A.wbs);
    if (ok)
        PP_CUDA_TRY(cudaMemsetAsync(ok, 1, n, stream)); // the encoder cannot fail (src/encode.c:17-145 returns true)
    {
        ProfScope prof(ctx, PROF_ENCODE, stream);
        rs255_lfsr_kernel<MODE_ENCODE><<<lfsr_grid(ctx, n), 64 * kPairs, kPairs * kSmemEncode, stream>>>(A);
    }
    ctx->launches++;
    return check(cudaGetLastError(), "rs255 encode launch");
}

int rs255_decode(poporon_cuda_rs *ctx, int mode, uint8_t *data, size_t data_stride, size_t size, uint8_t *parity,
                 size_t parity_stride, size_t n, const uint32_t *pos, size_t pos_stride, const uint32_t *cnt,
                 const uint16_t *extsyn, uint8_t *ok, uint32_t *corrected, cudaStream_t stream)
{
    if (n == 0)
        return 0;
    const int nr = ctx->code.nroots; // 32, 16 or 8
    // (the kernel reads 32 slots of every position row, as quad-words where it can: rows narrower than that - allowed
    // for the codes with 16 and 8 roots - keep the older path)
    const bool era_kernel = mode == POPORON_CUDA_MODE_ERASURE && pos_stride >= 32 && !ctx->no_erasure_kernel;
    int rc = ensure_scratch(ctx, n, era_kernel ? (size_t)kEraWarps * kEraScratchWords : (size_t)kSolveWarpsPerSM * kSolveScratchWords);
    if (rc != 0)
        return rc;
    const size_t tiles = 4 * ((n + 4095) / 4096);
    uint32_t *dmask = ctx->d_queue + ((n + 3) & ~(size_t)3);
    PP_CUDA_TRY(cudaMemsetAsync(ctx->d_counters, 0, kCountersBytes, stream));
    LfsrArgs A{};
    A.stream = data;
    A.stride = data_stride;
    A.len = (int)size + kPar;
    A.n = n;
    A.gend = data + (n - 1) * data_stride + size + kPar;
    A.dmask = dmask;
    A.syn = ctx->d_rem;
    FftShape shape;
    if (mode == POPORON_CUDA_MODE_EXTSYN) {
        // the syndromes come from the caller: no screen pass, the rows and the queue are built from them
        ProfScope prof(ctx, PROF_COMPACT, stream);
        rs255_extsyn_ingest_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(
            nr, extsyn, n, ctx->d_tables, ctx->d_rem, ctx->d_queue, ctx->d_counters, ok, corrected);
    } else if ((!ctx->use_lfsr || ctx->code.nroots != 32) &&
               rs255_fft_shape_ok(ctx->code.nroots, 1, data, data_stride, size, parity, parity_stride, &shape)) {
        rc = ensure_slab(ctx);
        if (rc != 0)
            return rc;
        // the zero-syndrome exit rides in the screen kernel when the status arrays allow vector stores
        const bool fused = (((uintptr_t)ok | (uintptr_t)corrected) & 15u) == 0 && !ctx->no_fuse;
        if (fused)
            ctx->launches -= 1; // (no separate compaction kernel: see the count at the end)
        {
            ProfScope prof(ctx, PROF_SCREEN, stream);
            rc = rs255_fft_screen(ctx, shape, data, data_stride, size, parity, parity_stride, n, ctx->d_slab, dmask,
                                  ctx->d_rem, stream, ctx->d_queue, ctx->d_counters, ok, corrected, fused);
        }
        if (rc != 0)
            return rc;
        if (!fused) {
            ProfScope prof(ctx, PROF_COMPACT, stream);
            rc = rs255_fft_compact(ctx, dmask, n, ctx->d_queue, ctx->d_counters, ok, corrected, stream);
        }
        if (rc != 0)
            return rc;
    } else {
        uint64_t invperm;
        tile_geometry(data, data_stride, &A.perm, &invperm, &A.wbs);
        {
            ProfScope prof(ctx, PROF_SCREEN, stream);
            rs255_lfsr_kernel<MODE_SCREEN><<<lfsr_grid(ctx, n), 64 * kPairs, kPairs * kSmemScreen, stream>>>(A);
        }
        {
            ProfScope prof(ctx, PROF_COMPACT, stream);
            rs255_compact_kernel<<<(unsigned)((tiles * kTile + 255) / 256), 256, 0, stream>>>(
                dmask, n, invperm, ctx->d_queue, ctx->d_counters, ok, corrected);
        }
    }
    // one resident wave of the general stage (its shared-memory work arrays limit it to a few CTAs per SM); a grid
    // larger than one wave would leave a thin second wave behind the grid-stride loop
    const unsigned cblocks = (unsigned)(ctx->sm_count * ctx->correct_ctas_per_sm);
    const uint32_t *cq = ctx->d_queue, *cc = ctx->d_counters;
    const uint32_t *era_counter = ctx->d_counters + 1; // rows for the literal stage
    {
        // bitsliced solver first (erasure mode: for the rows it can prove equal to plain decoding, see erasure_row_covers_roots) (with the caller's syndromes error_correction_u8 runs exactly as in plain mode); what it declines lands in queue2 for the literal stage below
        uint32_t *queue2 = dmask + tiles * 64;
        const size_t groups = (n + 1023) / 1024;
        const size_t cap = (size_t)ctx->sm_count * kSolveWarpsPerSM;
        uint32_t *scratch = reinterpret_cast<uint32_t *>(ctx->d_rem + tiles * kTile * 32);
        if (era_kernel) {
            // erasure mode: the erasure-only kernel takes every regular row whose errors sit on its declared positions.
            // The queue is first ordered by declared count (into the buffer of the second queue; what the kernel
            // declines then goes where the unordered queue was).
            {
                ProfScope prof(ctx, PROF_COMPACT, stream);
                const unsigned sblocks = (unsigned)ctx->sm_count * 8u;
                rs255_era_hist_kernel<<<sblocks, 256, 0, stream>>>(ctx->d_queue, ctx->d_counters, cnt, ctx->d_counters + kCntHist);
                rs255_era_scatter_kernel<<<sblocks, 256, 0, stream>>>(ctx->d_queue, ctx->d_counters, cnt, ctx->d_counters + kCntHist,
                                                                      ctx->d_counters + kCntCursor, queue2);
                ctx->launches += 2;
            }
            {
                ProfScope prof(ctx, PROF_ERASURE, stream);
                const size_t ecap = (size_t)ctx->sm_count * kEraWarps;
                const size_t warps = groups < ecap ? groups : ecap;
                rs255_erasure_kernel<<<(unsigned)((warps + kEraWarps - 1) / kEraWarps), 32 * kEraWarps, kEraSmem, stream>>>(
                    nr, size, queue2, ctx->d_counters, ctx->d_rem, scratch, ctx->d_queue, ctx->d_counters + 1, ok, corrected,
                    data, data_stride, pos, pos_stride, cnt);
            }
            // what it declined (now in the buffer of the first queue, counted in counters[1]), ordered by declared count
            // again, goes through the second bitsliced stage; what that declines ends up in the first buffer once more
            if (!ctx->no_ee_kernel && nr == 32) {
                uint32_t *sorted = queue2; // (the ordered queue of the erasure kernel is dead)
                {
                    ProfScope prof2(ctx, PROF_COMPACT, stream);
                    const unsigned sblocks = (unsigned)ctx->sm_count * 8u;
                    rs255_era_hist_kernel<<<sblocks, 256, 0, stream>>>(ctx->d_queue, ctx->d_counters + 1, cnt, ctx->d_counters + kCntHist2);
                    rs255_era_scatter_kernel<<<sblocks, 256, 0, stream>>>(ctx->d_queue, ctx->d_counters + 1, cnt, ctx->d_counters + kCntHist2,
                                                                          ctx->d_counters + kCntCursor2, sorted);
                    ctx->launches += 2;
                }
                ProfScope prof3(ctx, PROF_EE, stream);
                const size_t ewarps = groups < (size_t)ctx->sm_count * kEeWarps ? groups : (size_t)ctx->sm_count * kEeWarps;
                rs255_ee_kernel<<<(unsigned)((ewarps + kEeWarps - 1) / kEeWarps), 32 * kEeWarps, PP_EE_B_SMEM ? kEeWarps * 128 * 128 : 0, stream>>>(
                    size, sorted, ctx->d_counters + 1, ctx->d_rem, scratch, ctx->d_queue, ctx->d_counters + 2, ok, corrected, data,
                    data_stride, pos, pos_stride, cnt, ctx->ee_min_rows);
                ctx->launches += 1;
                era_counter = ctx->d_counters + 2;
            }
            queue2 = ctx->d_queue; // (the literal stage reads what was declined)
        } else {
            ProfScope prof(ctx, PROF_SOLVE, stream);
            const size_t warps = groups < cap ? groups : cap;
            rs255_solve_kernel<<<(unsigned)((warps + kSolveWarpsPerCTA - 1) / kSolveWarpsPerCTA), 32 * kSolveWarpsPerCTA,
                                 kSolveSmem, stream>>>(
                nr, size, ctx->d_queue, ctx->d_counters, ctx->d_rem, scratch, queue2, ctx->d_counters + 1, ok, corrected,
                data, data_stride, parity, parity_stride, mode == POPORON_CUDA_MODE_ERASURE ? pos : nullptr, pos_stride,
                mode == POPORON_CUDA_MODE_ERASURE ? cnt : nullptr);
        }
        cq = queue2;
        cc = era_counter;
        ctx->launches += 1;
    }
    {
        ProfScope prof(ctx, PROF_GENERAL, stream);
        rs255_correct_kernel<<<cblocks, kWorkThreads, kCorrectSmem, stream>>>(ctx->code, ctx->d_tables, mode, data, data_stride, size,
                                                          parity, parity_stride, cq, cc, ctx->d_rem, pos, pos_stride,
                                                          cnt, ok, corrected);
    }
    ctx->launches += 3;
    return check(cudaGetLastError(), "rs255 decode launch");
}

int rs255_syndromes(poporon_cuda_rs *ctx, const uint8_t *data, size_t data_stride, size_t size, const uint8_t *parity,
                    size_t parity_stride, size_t n, uint16_t *out, cudaStream_t stream)
{
    if (n == 0)
        return 0;
    FftShape shape;
    if (!rs255_fft_shape_ok(ctx->code.nroots, 1, data, data_stride, size, parity, parity_stride, &shape))
        return POPORON_CUDA_ERR_UNSUPP;
    int rc = ensure_scratch(ctx, n, (size_t)kSolveWarpsPerSM * kSolveScratchWords);
    if (rc == 0)
        rc = ensure_slab(ctx);
    if (rc != 0)
        return rc;
    uint32_t *dmask = ctx->d_queue + ((n + 3) & ~(size_t)3);
    {
        ProfScope prof(ctx, PROF_SCREEN, stream);
        rc = rs255_fft_screen(ctx, shape, data, data_stride, size, parity, parity_stride, n, ctx->d_slab, dmask,
                              ctx->d_rem, stream);
    }
    if (rc != 0)
        return rc;
    const int nr = ctx->code.nroots;
    rs255_syn_export_kernel<<<(unsigned)((n * (size_t)(nr / 4) + 255) / 256), 256, 0, stream>>>(nr, dmask, ctx->d_rem, n,
                                                                                               ctx->d_tables, out);
    ctx->launches += 2;
    return check(cudaGetLastError(), "rs255 syndrome export launch");
}

} // namespace ppcu
