// rs255_fft.cu — syndromes (and through them the parity) of RS(255,223) as a 15 x 17 Cooley-Tukey transform on
// bit-planes.  sm_100a only.
//
// What replaces what (reference tree):
//   calculate_syndrome_u8 (src/decode.c:232-416) ... rs255_fft_kernel<SCREEN>: "is the codeword clean" + S_1..S_32
//   rs_encode (src/encode.c:17-145) ................ rs255_fft_kernel<ENCODE>: parity = V^-1 * syndromes(message * x^32)
//
// The arithmetic (tools/gen_rs255_fft.py derives, checks and writes the XOR networks, rs255_fft_gen.cuh):
//   e = 17 d + c,  R_c[k] = sum_d r_(17d+c) beta^(dk)  (15-point DFT, twiddles in GF(16), 488 lane-ops per class)
//   S_j = sum_c alpha^(jc) R_c[j mod 15]               (Horner in c, 486 lane-ops per class)
// = 18.6 K lane-ops per 32 codewords against 80.6 K for 32 Horner / LFSR recurrences over 255 symbols.
//
// A transform needs a whole codeword at once, and 1024 codewords in bit-plane form (one plane word = 32
// codewords, 32 lanes) are 255 KB: more than a CTA's shared memory.  So a CTA works on a tile of 1024 consecutive
// codewords in two phases, with the planes parked in a CTA-private 255 KB slab of global memory that is written
// and read back by the same SM and therefore lives in L2:
//   phase A  every warp takes sub-tiles of 32 consecutive codewords (one lane-slot of the tile).  The rows arrive
//            by ONE bulk copy (cp.async.bulk + mbarrier: the 32 rows are contiguous in memory) into the warp's
//            buffer, which is refilled as soon as the warp has read it.  Lane x transposes word-column x of the 32
//            rows (4 symbols x 32 codewords) into bit-planes, changes basis and stores each symbol's 8 planes
//            (32 bytes) to slab[slot(e)][lane-slot].
//   phase B  warps 0-7 ("S1") run the 15-point DFTs, one GF(16) half of sixteen lane-slots per task (class c = 15
//            slab slots = 15 KB contiguous), and hand R_c to warps 8-15 ("S2") through a 4-slot shared-memory ring
//            (mbarrier full/empty); S2 warp u owns four syndromes fed by two residues j mod 15 and runs their Horner
//            recurrences in registers.
//   output   SCREEN: a dirty bit per codeword + the 32 syndrome bytes of dirty codewords.  ENCODE: the syndrome planes
//            of message * x^32 go to global memory; rs255_parity_kernel (V^-1) and rs255_rows_kernel finish the job.
#include "rs255_fft.cuh"

#include <stdlib.h>

#include <utility>

#include "rs255_fft_gen.cuh"

namespace ppcu {

namespace {

constexpr int kTile = 1024;  // codewords per CTA tile
// One CTA of 16 warps per SM (128 registers per thread, four warps per scheduler), all warps in the same phase of the
// same tile.  Measured alternatives (16 Mi codewords, dirty screen / clean decode): 8 warps at 208 registers 2.61 /
// 2.20 ms; this 2.47 / 2.08 ms; 16 warps as a software pipeline of specialised warps (8 byte->plane, 4 + 4 transform)
// coupled by mbarriers over two slabs 3.5 / 2.47 ms - the byte->plane warps became the critical path; two independent
// CTAs of 8 warps per SM 2.61 / 2.18 ms.
constexpr int kWarps = 16;
constexpr int kS1Warps = 8, kS2Warps = 8;            // phase B roles: warps 0..7 / 8..15
constexpr int kSubPerWarp = 32 / kWarps;
constexpr int kSlotBytes = 1024;                    // one symbol of the tile: [lane-slot 32][8 planes]
constexpr size_t kSlabBytes = 255 * kSlotBytes;     // per CTA
constexpr int kRingSlots = 4;
constexpr int kRingSlotBytes = 15 * 1024;           // R_c: [k 15][half 2][lane 32][4 planes]
constexpr int kRingBytes = kRingSlots * kRingSlotBytes;
constexpr int kMaxBuf = 8448;                       // bytes of one sub-tile buffer (rows of 32 codewords)
constexpr int kBarBytes = 2048;                      // mbarriers [0, 192) + the S2 warps' dirty words [256, 1280)
constexpr int kOutStageBytes = 256 * 128;            // SCREEN: the 256 syndrome planes of a tile, [plane][lane-slot]

enum { MODE_SCREEN = 0, MODE_ENCODE = 1 };

struct FftArgs {
    const uint8_t *data;   // message rows
    size_t dstride;
    int size;              // message bytes per codeword, 1..255 - npar
    int npar;              // parity bytes per codeword = roots of the code: 32, 16 or 8 (RS(255,223) / (255,239) / (255,247))
    const uint8_t *parity; // SCREEN: parity rows (32 bytes each)
    size_t pstride;
    int packed;            // SCREEN: parity rows interleaved with the message rows (one stride, one span per sub-tile)
    int prow_off;          // packed: parity - data (>= size; == size for packed records)
    size_t n;
    uint8_t *slab;         // gridDim.x slabs
    int bufbytes;          // bytes of one sub-tile buffer
    int poff;              // offset of the parity span inside a buffer (split layout)
    // SCREEN outputs
    uint32_t *dirty;       // bitmap, bit (cw & 31) of word cw >> 5
    uint8_t *syn;          // [tiles * 1024][32]: rows of the tiles that hold a dirty codeword
    // SCREEN, fused zero-syndrome exit (when ok / corrected are 16-byte aligned; else rs255_fft_compact_kernel runs):
    uint32_t *queue, *counter; // dirty codeword indices | 3 << 30, their count
    uint8_t *ok;               // nullable; clean: 1, dirty: 0 for now (the correction stages write the outcome)
    uint32_t *corrected;       // nullable; 0
    int fused;
    // ENCODE output: the syndromes of message * x^32 as tower planes, [tile][j - 1][lane-slot][8 planes] words
    // (rs255_parity_kernel turns them into parity rows)
    uint32_t *synplanes;
};

// ---------------------------------------------------------------------------------------------------
// small device helpers
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ uint32_t lds32(uint32_t a)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts32(uint32_t a, uint32_t v)
{
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ void lds128(uint32_t a, uint32_t (&v)[4])
{
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(a));
}
__device__ __forceinline__ void sts128(uint32_t a, const uint32_t (&v)[4])
{
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]) : "memory");
}
__device__ __forceinline__ void ldg256(const void *p, uint32_t (&lo)[4], uint32_t (&hi)[4])
{
    asm volatile("ld.global.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(lo[0]), "=r"(lo[1]), "=r"(lo[2]), "=r"(lo[3]), "=r"(hi[0]), "=r"(hi[1]), "=r"(hi[2]), "=r"(hi[3])
                 : "l"(p)
                 : "memory");
}
__device__ __forceinline__ void stg256(void *p, const uint32_t (&v)[8])
{
    asm volatile("st.global.v8.b32 [%8], {%0, %1, %2, %3, %4, %5, %6, %7};" ::"r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]),
                 "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "l"(p)
                 : "memory");
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.shared::cta.b64 st, [%0];\n}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n}" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        " .reg .pred p;\n"
        "WAIT_%=:\n"
        " mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        " @p bra DONE_%=;\n"
        " bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}" ::"r"(bar),
        "r"(parity)
        : "memory");
}
// one contiguous span global -> shared memory (both 16-byte aligned, bytes a multiple of 16), completion on an mbarrier
__device__ __forceinline__ void bulk_load(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}

// (a & m) | (b & ~m) in one LOP3
__device__ __forceinline__ uint32_t merge_bits(uint32_t a, uint32_t b, uint32_t m)
{
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xE4;" : "=r"(d) : "r"(a), "r"(b), "r"(m));
    return d;
}
__device__ __forceinline__ uint32_t xor3(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// 32x32 bit-matrix transpose in registers: x[r] bit c <-> x[c] bit r (an involution)
__device__ __forceinline__ void transpose32(uint32_t (&x)[32])
{
#pragma unroll
    for (int r = 0; r < 16; r++) {
        const uint32_t a = x[r], b = x[r + 16];
        x[r] = __byte_perm(a, b, 0x5410);
        x[r + 16] = __byte_perm(a, b, 0x7632);
    }
#pragma unroll
    for (int r = 0; r < 32; r++) {
        if (r & 8)
            continue;
        const uint32_t a = x[r], b = x[r + 8];
        x[r] = __byte_perm(a, b, 0x6240);
        x[r + 8] = __byte_perm(a, b, 0x7351);
    }
#pragma unroll
    for (int r = 0; r < 32; r++) {
        if (r & 4)
            continue;
        const uint32_t a = x[r], b = x[r + 4];
        x[r] = merge_bits(a, b << 4, 0x0F0F0F0Fu);
        x[r + 4] = merge_bits(a >> 4, b, 0x0F0F0F0Fu);
    }
#pragma unroll
    for (int r = 0; r < 32; r++) {
        if (r & 2)
            continue;
        const uint32_t a = x[r], b = x[r + 2];
        x[r] = merge_bits(a, b << 2, 0x33333333u);
        x[r + 2] = merge_bits(a >> 2, b, 0x33333333u);
    }
#pragma unroll
    for (int r = 0; r < 32; r++) {
        if (r & 1)
            continue;
        const uint32_t a = x[r], b = x[r + 1];
        x[r] = merge_bits(a, b << 1, 0x55555555u);
        x[r + 1] = merge_bits(a >> 1, b, 0x55555555u);
    }
}

// slab slot of the symbol with exponent e (< 255): class-major, slot = (e mod 17) * 15 + e / 17
__device__ __forceinline__ uint32_t slot_of(uint32_t e)
{
    const uint32_t d = (e * 241u) >> 12; // e / 17 for e < 255
    return (e - 17u * d) * 15u + d;
}

// all 16 subset-XORs of four planes
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

// stage word of (lane-slot ls, codeword i of the slot, word w of its row): rows of 8 words; the row of (ls, i) in
// slot ls * 32 + sigma with sigma mod 4 = ((i >> 2) + (ls >> 3)) mod 4, word w at position (w ^ ls) & 7.  Bank =
// 8 * (sigma mod 4) + position: the 32 lanes of a store (fixed i, w) hit 32 different banks, and so do the loads of
// the rows i, i + 4, i + 8, i + 12 of one lane-slot (the four rows of a copy-out step).
__device__ __forceinline__ uint32_t row_word(uint32_t ls, uint32_t i, uint32_t w)
{
    const uint32_t sigma = (((i >> 2) + (ls >> 3)) & 3u) | ((i & 3u) << 2) | (i & 16u);
    return ((ls * 32u + sigma) * 8u + ((w ^ ls) & 7u)) * 4u;
}

// ---------------------------------------------------------------------------------------------------
// shared memory map
//   [0, kBarBytes)                       mbarriers: buffer-full [kWarps], ring full [4], ring empty [4]; nz words
//   [kBarBytes, +kRingBytes)             S1 -> S2 ring
//   [.., + kWarps * bufbytes)            sub-tile buffers, one per warp
//   [.., + kOutStageBytes)               SCREEN: the output stage (256 planes x 128 B)
// ---------------------------------------------------------------------------------------------------
struct Smem {
    uint32_t base, bars, ring, bufs, outstage;
    int bufbytes;
    __device__ __forceinline__ uint32_t bar_buf(int w) const { return bars + 8u * (uint32_t)w; }
    __device__ __forceinline__ uint32_t bar_full(int s) const { return bars + 128u + 8u * (uint32_t)s; }
    __device__ __forceinline__ uint32_t bar_empty(int s) const { return bars + 160u + 8u * (uint32_t)s; }
    __device__ __forceinline__ uint32_t nzword(int u, int lane) const { return bars + 256u + 4u * (uint32_t)(32 * u + lane); }
    __device__ __forceinline__ uint32_t buf(int w) const { return bufs + (uint32_t)(w * bufbytes); }
    // plane p (0..255) of the output stage, this lane
    __device__ __forceinline__ uint32_t outplane(int p, int lane) const { return outstage + (uint32_t)(p * 128 + lane * 4); }
};

// ---------------------------------------------------------------------------------------------------
// phase A
// ---------------------------------------------------------------------------------------------------
// Issue the copy of the sub-tile that starts at codeword cwb (rows <= 32 of them exist) into buffer `bufaddr`.
// The data span starts (g & 15) bytes into the buffer.  Called by the whole warp (lane 0 issues the bulk copies,
// the other lanes move the few bytes that are not a multiple of 16).
template <int MODE>
__device__ __forceinline__ void issue_subtile(const FftArgs &A, uint32_t bufaddr, uint32_t bar, size_t cwb, int rows, int lane)
{
    __syncwarp(); // every lane is done reading the buffer's previous contents
    uint32_t tx = 0;
    const int rowlen = (MODE == MODE_SCREEN && A.packed) ? A.prow_off + A.npar : A.size;
    const uint8_t *g = A.data + cwb * A.dstride;
    const uint32_t a = (uint32_t)((uintptr_t)g & 15u);
    const uint32_t total = rows > 0 ? a + (uint32_t)(rows - 1) * (uint32_t)A.dstride + (uint32_t)rowlen : 0u;
    const uint32_t bulk = total & ~15u;
    tx += bulk;
    uint32_t bulk_p = 0, total_p = 0, ap = 0;
    const uint8_t *gp = nullptr;
    if (MODE == MODE_SCREEN && !A.packed) {
        gp = A.parity + cwb * A.pstride;
        ap = (uint32_t)((uintptr_t)gp & 15u);
        total_p = rows > 0 ? ap + (uint32_t)(rows - 1) * (uint32_t)A.pstride + (uint32_t)A.npar : 0u;
        bulk_p = total_p & ~15u;
        tx += bulk_p;
    }
    if (lane == 0) {
        mbar_expect_tx(bar, tx);
        if (bulk)
            bulk_load(bufaddr, g - a, bulk, bar);
        if (bulk_p)
            bulk_load(bufaddr + (uint32_t)A.poff, gp - ap, bulk_p, bar);
    }
    // tails: plain byte copies (a bulk copy moves multiples of 16 bytes and must not read past the batch)
    if ((uint32_t)lane < (total & 15u)) {
        const uint8_t v = (g - a)[bulk + (uint32_t)lane];
        asm volatile("st.shared.u8 [%0], %1;" ::"r"(bufaddr + bulk + (uint32_t)lane), "r"((uint32_t)v) : "memory");
    }
    if (MODE == MODE_SCREEN && !A.packed && (uint32_t)lane < (total_p & 15u)) {
        const uint8_t v = (gp - ap)[bulk_p + (uint32_t)lane];
        asm volatile("st.shared.u8 [%0], %1;" ::"r"(bufaddr + (uint32_t)A.poff + bulk_p + (uint32_t)lane), "r"((uint32_t)v) : "memory");
    }
}

// Transpose the sub-tile in `bufaddr` (lane-slot `ls` of the tile) into the slab.
template <int MODE>
__device__ __forceinline__ void transpose_subtile(const FftArgs &A, uint32_t bufaddr, size_t cwb, int ls, uint8_t *slab, int lane)
{
    const int L = A.size + A.npar;
    const int ncd = (A.size + 3) >> 2;                                   // word-columns of the message
    const int ncols = MODE == MODE_SCREEN ? ncd + (A.npar >> 2) : ncd;    // + those of the parity
    const uint32_t ad = (uint32_t)((uintptr_t)(A.data + cwb * A.dstride) & 15u);
    uint32_t pbase, pstr;
    if (MODE == MODE_SCREEN) {
        if (A.packed) {
            pbase = bufaddr + ad + (uint32_t)A.prow_off;
            pstr = (uint32_t)A.dstride;
        } else {
            pbase = bufaddr + (uint32_t)A.poff + (uint32_t)((uintptr_t)(A.parity + cwb * A.pstride) & 15u);
            pstr = (uint32_t)A.pstride;
        }
    } else {
        pbase = bufaddr;
        pstr = 0;
    }
#pragma unroll 1
    for (int col0 = 0; col0 < ncols; col0 += 32) {
        const int col = col0 + lane;
        uint32_t base, stride;
        int nvalid, e0;
        if (col < ncd) {
            base = bufaddr + ad + 4u * (uint32_t)col;
            stride = (uint32_t)A.dstride;
            nvalid = A.size - 4 * col;
            e0 = L - 1 - 4 * col;
        } else if (col < ncols) {
            const int pc = col - ncd;
            base = pbase + 4u * (uint32_t)(pc & 7);
            stride = pstr;
            nvalid = 4;
            e0 = A.npar - 1 - 4 * (pc & 7);
        } else {
            // idle lane of the last trip: it re-reads the words of the last message column (a broadcast, not a second
            // wavefront on some other bank) and stores nothing
            base = bufaddr + ad + 4u * (uint32_t)(ncd - 1);
            stride = (uint32_t)A.dstride;
            nvalid = 0;
            e0 = 0;
        }
        // rows i and i + 4 are 4 * stride bytes apart: same byte offset inside their words.  Four aligned row bases
        // and four PRMT selectors serve all 32 rows.
        uint32_t rb[4], sel[4];
#pragma unroll
        for (int v = 0; v < 4; v++) {
            const uint32_t addr = base + (uint32_t)v * stride;
            rb[v] = addr & ~3u;
            sel[v] = 0x3210u + 0x1111u * (addr & 3u);
        }
        const uint32_t s4 = 4u * stride;
        uint32_t X[32];
#pragma unroll
        for (int i = 0; i < 32; i++) {
            const uint32_t addr = rb[i & 3] + (uint32_t)(i >> 2) * s4;
            const uint32_t lo = lds32(addr), hi = lds32(addr + 4u);
            X[i] = __byte_perm(lo, hi, sel[i & 3]);
        }
        transpose32(X); // X[8 y + t] = plane t of byte y of the column
        // basis change in place, then the four plane stores together: their source registers are not written
        // again before the next column's loads have come back (a store that waits in the LSU queue holds its
        // registers: rewriting them at once cost 20 % of this loop's time in scoreboard stalls)
#pragma unroll
        for (int y = 0; y < 4; y++) {
            uint32_t(&xs)[8] = *reinterpret_cast<uint32_t(*)[8]>(&X[8 * y]);
            rsfft::conv_in(xs, xs);
        }
#pragma unroll
        for (int y = 0; y < 4; y++) {
            if (y < nvalid) {
                uint32_t(&xs)[8] = *reinterpret_cast<uint32_t(*)[8]>(&X[8 * y]);
                stg256(slab + (size_t)slot_of((uint32_t)(e0 - y)) * kSlotBytes + (size_t)ls * 32, xs);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// phase B
// ---------------------------------------------------------------------------------------------------
// S1: 34 half-class tasks g2 = 2 t + lh (class c = 16 - t, so that S2 meets the classes in Horner order; lh = which
// sixteen lane-slots of the tile).  A lane owns one GF(16) half (4 planes) of one lane-slot: lane = 16 * half + slot,
// so that a warp reads 512 contiguous bytes per symbol, a thread holds 15 x 4 planes only, and the ring stores of a
// quarter warp are 128 contiguous bytes (no bank conflicts).
__device__ __forceinline__ void ldg128(const void *p, uint32_t (&v)[4])
{
    asm volatile("ld.global.cg.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "l"(p) : "memory"); // (L2 only: the slab is rewritten every other tile)
}
__device__ __forceinline__ void s1_run(const Smem &sm, const uint8_t *slab, int u, int lane, uint32_t gtask)
{
    uint32_t x[15][4];
    const int half = lane >> 4, sl = lane & 15;
#pragma unroll 1
    for (int g2 = u; g2 < 34; g2 += kS1Warps) {
        const int t = g2 >> 1, lh = g2 & 1;
        const int c = 16 - t;
        const uint8_t *src = slab + (size_t)(c * 15) * kSlotBytes + (size_t)((lh * 16 + sl) * 32 + half * 16);
#pragma unroll
        for (int d = 0; d < 15; d++)
            ldg128(src + (size_t)d * kSlotBytes, x[d]);
        rsfft::dft15(x);
        const uint32_t g = gtask + (uint32_t)t;
        const int slot = (int)(g & (kRingSlots - 1));
        mbar_wait(sm.bar_empty(slot), ((g / kRingSlots) & 1u) ^ 1u);
        // ring slot: [k 15][half 2][lane-slot 32][4 planes]
        const uint32_t dst = sm.ring + (uint32_t)(slot * kRingSlotBytes + half * 512 + (lh * 16 + sl) * 16);
#pragma unroll
        for (int k = 0; k < 15; k++)
            sts128(dst + (uint32_t)(k * 1024), x[k]);
        __syncwarp();
        if (lane == 0)
            mbar_arrive(sm.bar_full(slot)); // (two arrivals complete a class: one per half of the lane-slots)
    }
}

// the syndromes of S2 warp U (0..7): four syndromes fed by two residues k = j mod 15.
//   U = 0: residues (1, 3): S_1, S_16, S_31, S_3        U = 1: residues (2, 3): S_2, S_17, S_32, S_18
//   U >= 2: residues (2U, 2U + 1 mod 15): S_2U, S_(2U+15), S_(2U+1), S_(2U+16)
__host__ __device__ constexpr int s2_k(int u, int i) { return u == 0 ? (i ? 3 : 1) : u == 1 ? (i ? 3 : 2) : (2 * u + i) % 15; }
__host__ __device__ constexpr int s2_kj(int u, int q) { return u < 2 ? (q < 3 ? 0 : 1) : q >> 1; } // residue feeding slot q
__host__ __device__ constexpr int s2_j(int u, int q)                                               // syndrome of slot q
{
    return u == 0 ? (q < 3 ? 1 + 15 * q : 3) : u == 1 ? (q < 3 ? 2 + 15 * q : 18) : (2 * u + (q >> 1)) + 15 * (q & 1);
}
constexpr bool s2_table_ok()
{
    bool seen[33] = {};
    for (int u = 0; u < 8; u++)
        for (int q = 0; q < 4; q++) {
            const int j = s2_j(u, q);
            if (j < 1 || j > 32 || seen[j] || j % 15 != s2_k(u, s2_kj(u, q)))
                return false;
            seen[j] = true;
        }
    return true;
}
static_assert(s2_table_ok(), "syndrome assignment");

template <int U, int... Q>
__device__ __forceinline__ void s2_step(uint32_t (&acc)[4][8], const uint32_t (&r)[2][8], std::integer_sequence<int, Q...>)
{
    (rsfft::hstep<s2_j(U, Q)>(acc[Q], r[s2_kj(U, Q)]), ...);
}

// S2: Horner over the 17 classes for this warp's four syndromes; leaves them (tower planes) in acc
template <int U>
__device__ __forceinline__ void s2_run(const Smem &sm, int lane, uint32_t gtask, uint32_t (&acc)[4][8])
{
#pragma unroll
    for (int q = 0; q < 4; q++)
#pragma unroll
        for (int t = 0; t < 8; t++)
            acc[q][t] = 0;
#pragma unroll 1
    for (int t = 0; t < 17; t++) {
        const uint32_t g = gtask + (uint32_t)t;
        const int slot = (int)(g & (kRingSlots - 1));
        mbar_wait(sm.bar_full(slot), (g / kRingSlots) & 1u);
        const uint32_t src = sm.ring + (uint32_t)(slot * kRingSlotBytes + lane * 16);
        uint32_t r[2][8];
#pragma unroll
        for (int i = 0; i < 2; i++) {
            uint32_t a[4], b[4];
            lds128(src + (uint32_t)(s2_k(U, i) * 1024), a);
            lds128(src + (uint32_t)(s2_k(U, i) * 1024 + 512), b);
#pragma unroll
            for (int q = 0; q < 4; q++) {
                r[i][q] = a[q];
                r[i][4 + q] = b[q];
            }
        }
        s2_step<U>(acc, r, std::make_integer_sequence<int, 4>{});
        __syncwarp(); // every lane has its operands in registers (and used them)
        if (lane == 0)
            mbar_arrive(sm.bar_empty(slot));
    }
}

// S2 epilogue.  SCREEN: dirty mask of the warp's syndromes + their standard-basis planes to the output stage.
// ENCODE: the tower planes go to the global syndrome-plane array of the tile (32 bytes per lane and syndrome).
template <int MODE, int U>
__device__ __forceinline__ uint32_t s2_store(const Smem &sm, int lane, uint32_t (&acc)[4][8], uint32_t *tileplanes, int npar)
{
    uint32_t nz = 0;
#pragma unroll
    for (int q = 0; q < 4; q++) {
        if (MODE == MODE_SCREEN) {
            uint32_t s[8];
            rsfft::conv_out(acc[q], s);
#pragma unroll
            for (int t = 0; t < 8; t++) {
                nz |= s2_j(U, q) <= npar ? s[t] : 0u; // (a 16- or 8-root code has no syndromes beyond its roots)
                sts32(sm.outplane(8 * (s2_j(U, q) - 1) + t, lane), s[t]);
            }
        } else {
            stg256(tileplanes + ((size_t)(s2_j(U, q) - 1) * 32 + (size_t)lane) * 8, acc[q]);
        }
    }
    return nz;
}

template <int MODE, int U>
__device__ __forceinline__ void s2_role(const Smem &sm, int lane, uint32_t gtask, uint32_t *tileplanes, int npar)
{
    uint32_t acc[4][8];
    s2_run<U>(sm, lane, gtask, acc);
    const uint32_t nz = s2_store<MODE, U>(sm, lane, acc, tileplanes, npar);
    if (MODE == MODE_SCREEN)
        sts32(sm.nzword(U, lane), nz);
}

// SCREEN output of a tile that holds a dirty codeword: its 1024 syndrome rows (32 bytes each, contiguous in syn).
// Word-column w is transposed by warp w < 8, staged in the (now idle) output stage and written as whole rows by all
// warps.  Called by the whole CTA; kept out of line so that its 32 transpose registers do not weigh on the register
// allocation of the main loop (the kernel sits at its 128-register limit).
__device__ __noinline__ void write_syndrome_rows(const Smem &sm, uint8_t *syn, size_t tile, int warp, int lane)
{
    uint32_t X[32];
    if (warp < 8) {
#pragma unroll
        for (int p = 0; p < 32; p++)
            X[p] = lds32(sm.outplane(32 * warp + p, lane));
        transpose32(X);
    }
    __syncthreads(); // every plane has been read
    if (warp < 8) {
#pragma unroll
        for (int i = 0; i < 32; i++)
            sts32(sm.outstage + row_word((uint32_t)lane, (uint32_t)i, (uint32_t)warp), X[i]);
    }
    __syncthreads();
    uint32_t *dst = reinterpret_cast<uint32_t *>(syn) + tile * (size_t)(kTile * 8);
    // a warp moves four rows i, i + 4, i + 8, i + 12 of one lane-slot per step (the four bank groups of the stage), eight
    // lanes per row: whole 32-byte sectors, no bank conflicts
    const uint32_t rr = (uint32_t)lane >> 3, k = (uint32_t)lane & 7u;
#pragma unroll 4
    for (uint32_t it = (uint32_t)warp; it < 256u; it += kWarps) {
        const uint32_t ls = it >> 3, q = it & 7u;
        const uint32_t i = (q & 3u) + 16u * (q >> 2) + 4u * rr;
        dst[(ls * 32u + i) * 8u + k] = lds32(sm.outstage + row_word(ls, i, k));
    }
}

template <int MODE>
__global__ void __launch_bounds__(32 * kWarps, 1) rs255_fft_kernel(FftArgs A)
{
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    Smem sm;
    sm.base = smem_u32(smem_raw);
    sm.bars = sm.base;
    sm.ring = sm.base + kBarBytes;
    sm.bufs = sm.ring + kRingBytes;
    sm.bufbytes = A.bufbytes;
    sm.outstage = sm.bufs + (uint32_t)(kWarps * A.bufbytes);
    uint8_t *slab = A.slab + (size_t)blockIdx.x * kSlabBytes;
    const size_t ntiles = (A.n + kTile - 1) / kTile;

    if (threadIdx.x == 0) {
        for (int w = 0; w < kWarps; w++)
            mbar_init(sm.bar_buf(w), 1);
        for (int s = 0; s < kRingSlots; s++) {
            mbar_init(sm.bar_full(s), 2);          // the two half-class tasks of S1
            mbar_init(sm.bar_empty(s), kS2Warps);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // symbols nobody writes are zero: exponents >= size + 32 (shortened code) and, for the encoder, the parity
    // positions e < 32
    {
        const int L = A.size + A.npar;
        const uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int e = warp; e < 255; e += kWarps)
            if (e >= L || (MODE == MODE_ENCODE && e < A.npar))
                stg256(slab + (size_t)slot_of((uint32_t)e) * kSlotBytes + (size_t)lane * 32, z);
    }
    __syncthreads();

    // sub-tile s of a tile is lane-slot ls = kWarps * s + warp; every warp has ONE buffer, refilled as soon as the warp
    // has read it (with four warps per scheduler the copy hides behind the other warps' transposes)
    uint32_t use = 0;             // completed uses of this warp's buffer (mbarrier phase)
    uint32_t gtask = 0;           // class tasks handed through the ring so far (ring phase)
    auto rows_of = [&](size_t tile, int ls) -> int {
        const size_t cwb = tile * kTile + (size_t)ls * 32;
        if (cwb >= A.n)
            return 0;
        const size_t left = A.n - cwb;
        return left < 32 ? (int)left : 32;
    };
    size_t tile = blockIdx.x;
    if (tile < ntiles)
        issue_subtile<MODE>(A, sm.buf(warp), sm.bar_buf(warp), tile * kTile + (size_t)warp * 32, rows_of(tile, warp), lane);

    for (; tile < ntiles; tile += gridDim.x) {
        // ---------------- phase A ----------------
#pragma unroll 1
        for (int s = 0; s < kSubPerWarp; s++) {
            const int ls = kWarps * s + warp;
            mbar_wait(sm.bar_buf(warp), use & 1u);
            use++;
            __syncwarp();
            transpose_subtile<MODE>(A, sm.buf(warp), tile * kTile + (size_t)ls * 32, ls, slab, lane);
            // next sub-tile (of this tile, or the first one of this CTA's next tile) into the same buffer
            if (s + 1 < kSubPerWarp) {
                const int nls = ls + kWarps;
                issue_subtile<MODE>(A, sm.buf(warp), sm.bar_buf(warp), tile * kTile + (size_t)nls * 32, rows_of(tile, nls), lane);
            } else if (tile + gridDim.x < ntiles) {
                const size_t nt = tile + gridDim.x;
                issue_subtile<MODE>(A, sm.buf(warp), sm.bar_buf(warp), nt * kTile + (size_t)warp * 32, rows_of(nt, warp), lane);
            }
        }
        __syncthreads(); // the slab is complete

        // ---------------- phase B ----------------
        uint32_t *tileplanes = MODE == MODE_ENCODE ? A.synplanes + tile * (size_t)(32 * 32 * 8) : nullptr; // kTilePlaneWords
        if (warp < kS1Warps) {
            s1_run(sm, slab, warp, lane, gtask);
        } else {
            switch (warp - kS1Warps) {
            case 0: s2_role<MODE, 0>(sm, lane, gtask, tileplanes, A.npar); break;
            case 1: s2_role<MODE, 1>(sm, lane, gtask, tileplanes, A.npar); break;
            case 2: s2_role<MODE, 2>(sm, lane, gtask, tileplanes, A.npar); break;
            case 3: s2_role<MODE, 3>(sm, lane, gtask, tileplanes, A.npar); break;
            case 4: s2_role<MODE, 4>(sm, lane, gtask, tileplanes, A.npar); break;
            case 5: s2_role<MODE, 5>(sm, lane, gtask, tileplanes, A.npar); break;
            case 6: s2_role<MODE, 6>(sm, lane, gtask, tileplanes, A.npar); break;
            default: s2_role<MODE, 7>(sm, lane, gtask, tileplanes, A.npar); break;
            }
        }
        gtask += 17;
        __syncthreads(); // S1 is done with the slab; SCREEN: all 256 syndrome planes are in the output stage

        // ---------------- output (SCREEN; the encoder's syndromes went to global memory) ----------------
        if (MODE == MODE_SCREEN) {
            const size_t cw0 = tile * kTile + (size_t)lane * 32; // this lane-slot's first codeword
            uint32_t valid = 0;                                  // bit i: codeword cw0 + i exists
            if (cw0 < A.n) {
                const size_t left = A.n - cw0;
                valid = left >= 32 ? 0xFFFFFFFFu : ((1u << (unsigned)left) - 1u);
            }
            uint32_t dirty = 0;
#pragma unroll
            for (int u = 0; u < kS2Warps; u++)
                dirty |= lds32(sm.nzword(u, lane));
            dirty &= valid;
            const bool any = __any_sync(0xFFFFFFFFu, dirty != 0); // (the same in every warp: lanes = the tile's lane-slots)
            // zero-syndrome exit: clean codewords are finished here (status stores spread over the CTA: one uint4 per
            // thread), dirty ones queued for the solver by warp 8
            const bool full_tile = tile * kTile + kTile <= A.n;
            if (A.fused && full_tile) {
                if (A.corrected && threadIdx.x < 256)
                    reinterpret_cast<uint4 *>(A.corrected + tile * kTile)[threadIdx.x] = make_uint4(0u, 0u, 0u, 0u);
                if (A.ok && threadIdx.x >= 256 && threadIdx.x < 320) {
                    const uint32_t q = threadIdx.x - 256u; // sixteen codewords: half q & 1 of lane-slot q >> 1
                    uint32_t d = 0;
#pragma unroll
                    for (int u = 0; u < kS2Warps; u++)
                        d |= lds32(sm.nzword(u, (int)(q >> 1)));
                    const uint32_t c = ~d >> (16u * (q & 1u));
                    uint32_t w[4];
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const uint32_t cj = c >> (4 * j);
                        w[j] = (cj & 1u) | ((cj & 2u) << 7) | ((cj & 4u) << 14) | ((cj & 8u) << 21);
                    }
                    reinterpret_cast<uint4 *>(A.ok + tile * kTile)[q] = make_uint4(w[0], w[1], w[2], w[3]);
                }
            }
            if (warp == 8) {
                if (cw0 < A.n)
                    A.dirty[cw0 >> 5] = dirty;
                if (A.fused) {
                    if (!full_tile) {
                        for (int i = 0; i < 32; i++)
                            if ((valid >> i) & 1u) {
                                if (A.ok)
                                    A.ok[cw0 + i] = (uint8_t)((~dirty >> i) & 1u);
                                if (A.corrected)
                                    A.corrected[cw0 + i] = 0u;
                            }
                    }
                    if (any) {
                        const uint32_t cnt = (uint32_t)__popc(dirty);
                        uint32_t pre = cnt; // inclusive prefix sum over the lanes
#pragma unroll
                        for (int d = 1; d < 32; d <<= 1) {
                            const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, pre, d);
                            if (lane >= d)
                                pre += v;
                        }
                        uint32_t base = 0;
                        if (lane == 31)
                            base = atomicAdd(A.counter, pre);
                        base = __shfl_sync(0xFFFFFFFFu, base, 31) + pre - cnt;
                        uint32_t bits = dirty;
                        while (bits) {
                            const int i = __ffs((int)bits) - 1;
                            bits &= bits - 1;
                            A.queue[base++] = (uint32_t)(cw0 + (size_t)i) | (3u << 30); // both syndrome halves are live
                        }
                    }
                }
            }
            if (any)
                write_syndrome_rows(sm, A.syn, tile, warp, lane);
            __syncthreads(); // the output stage and the dirty words are free again
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// The tail of the encoder (replaces the rest of rs_encode, src/encode.c:17-145: the reference divides by g(x);
// here parity = V^-1 S with V[j][t] = alpha^(j t)) in two kernels:
//
// rs255_parity_kernel   the dense 256 x 256 bit matrix V^-1 applied to the syndrome planes by four-Russians tables
//     with compile-time masks (rsfft::EncM).  The product is ~12 K straight-line instructions (190 KB of code), far
//     beyond the instruction cache, and cold code is fetched at about one 128-byte line per 150 cycles whatever
//     the number of warps (measured: 0.43 warp-instructions per clock and SM with 8 warps, profiles/r02_*): every
//     fetched instruction must serve as many warps as possible.  So a warp owns one tile (its 32 lanes = the tile's
//     32 lane-slots), the SIXTEEN warps of a CTA run the same stream on sixteen tiles (128 registers each: four
//     passes of 64 output planes) and meet at a barrier every four syndromes.  Output: parity bit-planes,
//     [tile][plane / 4][lane-slot][plane % 4] words, plane = 8 * byte + bit (byte in memory order).
// rs255_rows_kernel     planes -> bytes (32 x 32 bit transposes), staged per tile in shared memory as 32-byte rows
//     and written to the parity rows, which sit at arbitrary alignment inside packed records: four rows of the same
//     alignment class per step, one aligned 32-bit word per lane (head and tail of a misaligned row as bytes).
// ---------------------------------------------------------------------------------------------------
constexpr int kParWarps = 16;
constexpr int kParPlanes = 64;                         // output planes per pass
// the mask table of a code with NR roots (32: RS(255,223); 16, 8: the reference's other documented codes)
template <int NR> struct EncOf;
template <> struct EncOf<32> { using M = rsfft::EncM; };
template <> struct EncOf<16> { using M = rsfft::EncM16; };
template <> struct EncOf<8> { using M = rsfft::EncM8; };
constexpr size_t kTilePlaneWords = 32 * 32 * 8;        // one tile of syndrome planes / parity planes

struct ParArgs {
    const uint32_t *synplanes; // [tile][32][32][8]
    uint32_t *parplanes;       // [tile][64][32][4]
    size_t ntiles;
};

template <int NR, int H, int J, int QU>
__device__ __forceinline__ void par_one(uint32_t (&acc)[kParPlanes], const uint32_t (&T1)[16], const uint32_t (&T2)[16])
{
    constexpr unsigned m = EncOf<NR>::M::v.m[J][(kParPlanes / 8) * H + (QU >> 3)][QU & 7];
    acc[QU] = xor3(acc[QU], T1[m & 15], T2[m >> 4]);
}
template <int NR, int H, int J, int... QU>
__device__ __forceinline__ void par_acc(uint32_t (&acc)[kParPlanes], const uint32_t (&T1)[16], const uint32_t (&T2)[16],
                                        std::integer_sequence<int, QU...>)
{
    (par_one<NR, H, J, QU>(acc, T1, T2), ...);
}
template <int NR, int H, int J>
__device__ __forceinline__ void par_j(const uint32_t *tp, int lane, uint32_t (&in)[8], uint32_t (&acc)[kParPlanes])
{
    if constexpr (J < NR) {
        uint32_t nx[8], T1[16], T2[16];
        if constexpr (J + 1 < NR) {
            uint32_t lo[4], hi[4];
            ldg256(tp + ((size_t)(J + 1) * 32 + (size_t)lane) * 8, lo, hi);
#pragma unroll
            for (int t = 0; t < 4; t++) {
                nx[t] = lo[t];
                nx[4 + t] = hi[t];
            }
        }
        subset_table(in[0], in[1], in[2], in[3], T1);
        subset_table(in[4], in[5], in[6], in[7], T2);
        par_acc<NR, H, J>(acc, T1, T2, std::make_integer_sequence<int, kParPlanes>{});
        if constexpr ((J & 3) == 3)
            __syncthreads(); // keeps the CTA's warps on the same stretch of code
        if constexpr (J + 1 < NR) {
#pragma unroll
            for (int t = 0; t < 8; t++)
                in[t] = nx[t];
            par_j<NR, H, J + 1>(tp, lane, in, acc);
        }
    }
}

template <int NR, int H>
__device__ __forceinline__ void parity_pass(const uint32_t *tp, uint32_t *op, bool active, int lane)
{
    uint32_t acc[kParPlanes];
#pragma unroll
    for (int p = 0; p < kParPlanes; p++)
        acc[p] = 0;
    {
        uint32_t in[8], lo[4], hi[4];
        ldg256(tp + (size_t)lane * 8, lo, hi);
#pragma unroll
        for (int t = 0; t < 4; t++) {
            in[t] = lo[t];
            in[4 + t] = hi[t];
        }
        par_j<NR, H, 0>(tp, lane, in, acc);
    }
    if (active) {
#pragma unroll
        for (int q = 0; q < kParPlanes / 4; q++)
            *reinterpret_cast<uint4 *>(op + ((size_t)((kParPlanes / 4) * H + q) * 32 + (size_t)lane) * 4) =
                make_uint4(acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]);
    }
}
template <int NR, int... H>
__device__ __forceinline__ void parity_passes(const uint32_t *tp, uint32_t *op, bool active, int lane,
                                              std::integer_sequence<int, H...>)
{
    (parity_pass<NR, H>(tp, op, active, lane), ...);
}

template <int NR>
__global__ void __launch_bounds__(32 * kParWarps, 1) rs255_parity_kernel(ParArgs A)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t per_round = (size_t)gridDim.x * kParWarps;
    const size_t rounds = (A.ntiles + per_round - 1) / per_round;
    for (size_t it = 0; it < rounds; it++) {
        size_t tile = it * per_round + (size_t)blockIdx.x * kParWarps + (size_t)warp;
        const bool active = tile < A.ntiles;
        if (!active)
            tile = A.ntiles - 1; // walks through the barriers on a tile that exists, stores nothing
        parity_passes<NR>(A.synplanes + tile * kTilePlaneWords, A.parplanes + tile * kTilePlaneWords, active, lane,
                          std::make_integer_sequence<int, 8 * NR / kParPlanes>{});
    }
}

#ifndef PP_ROWS_SPLIT
#define PP_ROWS_SPLIT 2
#endif
constexpr int kRowSplit = PP_ROWS_SPLIT;      // warps that share a tile
constexpr int kRowTeams = 7;                  // tiles in flight per CTA (one 32 KB stage each)
constexpr int kRowWarps = kRowTeams * kRowSplit;
constexpr int kRowAhead = kRowSplit == 1 ? 3 : 2; // word-columns of planes in flight per warp
__device__ __forceinline__ void team_sync(int team)
{
    if (kRowSplit == 1)
        __syncwarp();
    else
        asm volatile("bar.sync %0, %1;" ::"r"(1 + team), "n"(32 * kRowSplit) : "memory");
}
constexpr int kRowStageBytes = kTile * 32; // the 32-byte parity rows of one tile
constexpr int kRowSmem = kRowTeams * kRowStageBytes;

struct RowArgs {
    const uint32_t *parplanes;
    size_t n, ntiles;
    uint8_t *parity;
    size_t pstride;
    int bytewise; // rows at addresses that are not all multiples of four
};

template <int NR> // parity bytes per row: 32, 16 or 8
__global__ void __launch_bounds__(32 * kRowWarps, 1) rs255_rows_kernel(RowArgs A)
{
    constexpr int KW = NR / 4;  // words per row
    constexpr int RPS = 32 / KW; // rows per copy-out step
    constexpr int CPW = KW / kRowSplit > 0 ? KW / kRowSplit : 1; // word-columns a warp of the team transposes
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // kRowSplit warps share a tile (and its stage): each transposes its share of the word-columns, then copies out its
    // share of the rows; they meet at a named barrier of their own.  Twice the warps per SM for the same shared memory.
    const int team = warp / kRowSplit, part = warp % kRowSplit;
    const uint32_t stage = smem_u32(smem_raw) + (uint32_t)(team * kRowStageBytes);
    for (size_t tile = (size_t)blockIdx.x * kRowTeams + (size_t)team; tile < A.ntiles; tile += (size_t)gridDim.x * kRowTeams) {
        const uint32_t *pp = A.parplanes + tile * kTilePlaneWords;
        // planes -> bytes: word-column w = parity bytes 4 w .. 4 w + 3 of the lane-slot's 32 codewords.  The planes of
        // the next columns are in flight while column w is transposed (one transpose is ~0.5 us, a DRAM read under
        // load more than 1 us).
        const int w0 = KW >= kRowSplit ? part * CPW : 0, w1 = KW >= kRowSplit ? w0 + CPW : (part == 0 ? KW : 0);
        uint4 nx[kRowAhead][8];
#pragma unroll
        for (int d = 0; d < kRowAhead; d++)
#pragma unroll
            for (int q = 0; q < 8; q++)
                nx[d][q] = *reinterpret_cast<const uint4 *>(pp + ((size_t)(8 * (w0 + d < KW ? w0 + d : KW - 1) + q) * 32 + (size_t)lane) * 4);
#pragma unroll 1
        for (int w = w0; w < w1; w++) {
            uint32_t X[32];
#pragma unroll
            for (int q = 0; q < 8; q++) {
                X[4 * q] = nx[0][q].x;
                X[4 * q + 1] = nx[0][q].y;
                X[4 * q + 2] = nx[0][q].z;
                X[4 * q + 3] = nx[0][q].w;
            }
            const int wn = w + kRowAhead < KW ? w + kRowAhead : KW - 1;
#pragma unroll
            for (int q = 0; q < 8; q++) {
#pragma unroll
                for (int d = 0; d + 1 < kRowAhead; d++)
                    nx[d][q] = nx[d + 1][q];
                nx[kRowAhead - 1][q] = *reinterpret_cast<const uint4 *>(pp + ((size_t)(8 * wn + q) * 32 + (size_t)lane) * 4);
            }
            transpose32(X);
#pragma unroll
            for (int i = 0; i < 32; i++)
                sts32(stage + row_word((uint32_t)lane, (uint32_t)i, (uint32_t)w), X[i]);
        }
        team_sync(team);
        // rows out.  A parity row inside a packed record covers two 32-byte sectors partially; the L2 merges partial
        // writes (a fill from DRAM per sector), which costs 2-3x the time of aligned contiguous rows (tools/
        // enc_probe.py).  Tried and slower: merging in the SM by read-modify-write of the two-sector window (0.59 ms
        // per 4 Mi rows), the same with the windows saved by the transform kernel while the records were in the L2
        // (0.57 ms: sixteen lanes per row make the copy-out instruction-bound), prefetching the sectors into the L2
        // first (0.40 ms), head/tail bytes in a loop of their own (0.34 ms).
        // Rows of a class c = r mod 4 have the same alignment a = address mod 4 (r * pstride mod 4 depends on r mod 4
        // only), so a is uniform over a step.  Lane = (row rr of four consecutive rows of the class, word k of eight):
        // one aligned 32-bit word per lane, its bytes picked from (previous word, this word) by one PRMT; the head
        // of a misaligned row leaves from lane k = 0, its tail from lane k = 7, as bytes.
        if (A.bytewise) {
            // rows at odd addresses: one byte per lane, 32 / NR whole rows per store instruction, so that every sector
            // a row touches gets ONE partial write (0.262 against 0.282 ms per 4 Mi rows with word stores + edge bytes;
            // for 4-byte aligned rows the word stores below are faster, 0.158 against 0.258 ms)
            const size_t cwt2 = tile * kTile;
            const uint32_t left2 = A.n - cwt2 >= (size_t)kTile ? (uint32_t)kTile : (uint32_t)(A.n - cwt2);
            constexpr uint32_t RPI = 32 / NR; // rows per instruction
            const uint32_t rsub = (uint32_t)lane / NR, b = (uint32_t)lane % NR;
            constexpr uint32_t span = (uint32_t)kTile / kRowSplit;
#pragma unroll 1
            for (uint32_t r0 = (uint32_t)part * span; r0 < ((uint32_t)part + 1u) * span; r0 += 8 * RPI) {
                uint32_t v[8];
#pragma unroll
                for (uint32_t u = 0; u < 8; u++) {
                    const uint32_t r = r0 + u * RPI + rsub;
                    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v[u]) : "r"(stage + row_word(r >> 5, r & 31u, b >> 2) + (b & 3u)));
                }
#pragma unroll
                for (uint32_t u = 0; u < 8; u++) {
                    const uint32_t r = r0 + u * RPI + rsub;
                    if (r < left2)
                        A.parity[(cwt2 + r) * A.pstride + b] = (uint8_t)v[u];
                }
            }
            team_sync(team);
            continue;
        }
        const size_t cwt = tile * kTile;
        const uint32_t left = A.n - cwt >= (size_t)kTile ? (uint32_t)kTile : (uint32_t)(A.n - cwt);
        const uint32_t rr = (uint32_t)lane / KW, k = (uint32_t)lane % KW;
#pragma unroll 1
        for (uint32_t c = 0; c < 4; c++) {
            const uint32_t a = (uint32_t)(((uintptr_t)A.parity + (cwt + c) * A.pstride) & 3u);
            const uint32_t sel = 0x3210u + 0x1111u * (4u - a); // bytes 4 - a .. 7 - a of (previous word, this word)
            const bool inner = a == 0 || k != 0;
#pragma unroll 1
            for (uint32_t it0 = (uint32_t)part * (256 / RPS / kRowSplit); it0 < ((uint32_t)part + 1u) * (256 / RPS / kRowSplit); it0 += 8) {
                uint32_t cur[8], prv[8];
#pragma unroll
                for (uint32_t u = 0; u < 8; u++) { // all sixteen loads first, then the stores
                    const uint32_t r = 4u * ((uint32_t)RPS * (it0 + u) + rr) + c;
                    cur[u] = lds32(stage + row_word(r >> 5, r & 31u, k));
                    prv[u] = lds32(stage + row_word(r >> 5, r & 31u, (k + KW - 1u) % KW));
                }
#pragma unroll
                for (uint32_t u = 0; u < 8; u++) {
                    const uint32_t r = 4u * ((uint32_t)RPS * (it0 + u) + rr) + c;
                    if (r >= left)
                        continue;
                    uint8_t *row = A.parity + (cwt + r) * A.pstride;
                    if (inner)
                        *reinterpret_cast<uint32_t *>(row - a + 4 * k) = __byte_perm(prv[u], cur[u], sel);
                    if (a != 0 && k == 0) { // head: row bytes 0 .. 3 - a
                        if (a == 1) {
                            row[0] = (uint8_t)cur[u];
                            *reinterpret_cast<uint16_t *>(row + 1) = (uint16_t)(cur[u] >> 8);
                        } else if (a == 2) {
                            *reinterpret_cast<uint16_t *>(row) = (uint16_t)cur[u];
                        } else {
                            row[0] = (uint8_t)cur[u];
                        }
                    }
                    if (a != 0 && k == KW - 1) { // tail: row bytes NR - a .. NR - 1, the top a bytes of the last word
                        if (a == 1) {
                            row[NR - 1] = (uint8_t)(cur[u] >> 24);
                        } else if (a == 2) {
                            *reinterpret_cast<uint16_t *>(row + NR - 2) = (uint16_t)(cur[u] >> 16);
                        } else {
                            *reinterpret_cast<uint16_t *>(row + NR - 3) = (uint16_t)(cur[u] >> 8);
                            row[NR - 1] = (uint8_t)(cur[u] >> 24);
                        }
                    }
                }
            }
        }
        team_sync(team);
    }
}

__global__ void __launch_bounds__(256)
rs255_fft_compact_kernel(const uint32_t *__restrict__ dirty, size_t n, uint32_t *__restrict__ queue,
                         uint32_t *__restrict__ counter, uint8_t *__restrict__ ok, uint32_t *__restrict__ corrected)
{
    const size_t cw = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t flag = 0;
    if (cw < n) {
        flag = (dirty[cw >> 5] >> (cw & 31u)) & 1u;
        if (!flag) {
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
            queue[base + __popc(ballot & ((1u << lane) - 1u))] = (uint32_t)cw | (3u << 30); // both syndrome halves are live
    }
}

int smem_bytes(int bufbytes)
{
    return kBarBytes + kRingBytes + kWarps * bufbytes + kOutStageBytes + 64;
}

} // namespace

bool rs255_fft_shape_ok(int npar, int op, const uint8_t *data, size_t data_stride, size_t size, const uint8_t *parity,
                        size_t parity_stride, FftShape *shape)
{
    if ((npar != 32 && npar != 16 && npar != 8) || size < 1 || size > (size_t)(255 - npar) || data_stride < size)
        return false;
    // "packed": every parity row sits between its message row and the next one (packed 255-byte records are the
    // case parity == data + size); the sub-tile is then one span
    const bool packed = op == 1 && parity_stride == data_stride && parity >= data + size &&
                        (size_t)(parity - data) + (size_t)npar <= data_stride;
    const size_t rowlen = packed ? (size_t)(parity - data) + (size_t)npar : size;
    size_t dspan = 31 * data_stride + rowlen + 15 + 8; // + alignment of the span + the funnel's second word
    dspan = (dspan + 15) & ~(size_t)15;
    size_t total = dspan, poff = 0;
    if (op == 1 && !packed) {
        if (parity_stride < (size_t)npar)
            return false;
        poff = dspan;
        size_t pspan = 31 * parity_stride + (size_t)npar + 15 + 8;
        pspan = (pspan + 15) & ~(size_t)15;
        total += pspan;
    }
    if (total > (size_t)kMaxBuf)
        return false;
    if (shape) {
        shape->npar = npar;
        shape->packed = packed ? 1 : 0;
        shape->prow_off = packed ? (int)(parity - data) : 0;
        shape->bufbytes = (int)total;
        shape->poff = (int)poff;
    }
    return true;
}

size_t rs255_fft_slab_bytes(int sm_count)
{
    return (size_t)sm_count * kSlabBytes;
}

int rs255_fft_init(poporon_cuda_rs *ctx)
{
    (void)ctx;
    PP_CUDA_TRY(cudaFuncSetAttribute(rs255_fft_kernel<MODE_SCREEN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     smem_bytes(kMaxBuf)));
    PP_CUDA_TRY(cudaFuncSetAttribute(rs255_fft_kernel<MODE_ENCODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     smem_bytes(kMaxBuf)));
    PP_CUDA_TRY(cudaFuncSetAttribute(rs255_rows_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, kRowSmem));
    PP_CUDA_TRY(cudaFuncSetAttribute(rs255_rows_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, kRowSmem));
    PP_CUDA_TRY(cudaFuncSetAttribute(rs255_rows_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, kRowSmem));
    return 0;
}

static unsigned fft_grid(const poporon_cuda_rs *ctx, size_t n)
{
    const size_t tiles = (n + kTile - 1) / kTile;
    return (unsigned)(tiles < (size_t)ctx->sm_count ? tiles : (size_t)ctx->sm_count);
}

int rs255_fft_screen(poporon_cuda_rs *ctx, const FftShape &shape, const uint8_t *data, size_t data_stride, size_t size,
                     const uint8_t *parity, size_t parity_stride, size_t n, uint8_t *slab, uint32_t *dirty,
                     uint8_t *syn, cudaStream_t stream, uint32_t *queue, uint32_t *counter, uint8_t *ok,
                     uint32_t *corrected, bool fused)
{
    FftArgs A{};
    A.data = data;
    A.dstride = data_stride;
    A.size = (int)size;
    A.npar = shape.npar;
    A.parity = parity;
    A.pstride = parity_stride;
    A.packed = shape.packed;
    A.prow_off = shape.prow_off;
    A.n = n;
    A.slab = slab;
    A.bufbytes = shape.bufbytes;
    A.poff = shape.poff;
    A.dirty = dirty;
    A.syn = syn;
    A.queue = queue;
    A.counter = counter;
    A.ok = ok;
    A.corrected = corrected;
    A.fused = fused ? 1 : 0;
    rs255_fft_kernel<MODE_SCREEN><<<fft_grid(ctx, n), 32 * kWarps, smem_bytes(shape.bufbytes), stream>>>(A);
    return check(cudaGetLastError(), "rs255 fft screen launch");
}

int rs255_fft_compact(poporon_cuda_rs *ctx, const uint32_t *dirty, size_t n, uint32_t *queue, uint32_t *counter,
                      uint8_t *ok, uint32_t *corrected, cudaStream_t stream)
{
    (void)ctx;
    rs255_fft_compact_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(dirty, n, queue, counter, ok, corrected);
    return check(cudaGetLastError(), "rs255 fft compact launch");
}

size_t rs255_fft_synplane_bytes(size_t n)
{
    return 2 * ((n + kTile - 1) / kTile) * kTilePlaneWords * 4; // syndrome planes + parity planes
}

int rs255_fft_encode(poporon_cuda_rs *ctx, const FftShape &shape, const uint8_t *data, size_t data_stride, size_t size,
                     uint8_t *parity, size_t parity_stride, size_t n, uint8_t *slab, uint32_t *synplanes,
                     cudaStream_t stream)
{
    FftArgs A{};
    A.data = data;
    A.dstride = data_stride;
    A.size = (int)size;
    A.npar = shape.npar;
    A.packed = 0;
    A.n = n;
    A.slab = slab;
    A.bufbytes = shape.bufbytes;
    A.poff = shape.poff;
    A.synplanes = synplanes;
    int rc;
    {
        ProfScope prof(ctx, PROF_ENCODE, stream);
        rs255_fft_kernel<MODE_ENCODE><<<fft_grid(ctx, n), 32 * kWarps, smem_bytes(shape.bufbytes), stream>>>(A);
        rc = check(cudaGetLastError(), "rs255 fft encode launch");
    }
    if (rc != 0)
        return rc;
    const size_t ntiles = (n + kTile - 1) / kTile;
    uint32_t *parplanes = synplanes + ntiles * kTilePlaneWords;
    {
        ParArgs P{};
        P.synplanes = synplanes;
        P.parplanes = parplanes;
        P.ntiles = ntiles;
        const size_t ctas = (ntiles + kParWarps - 1) / kParWarps;
        const unsigned grid = (unsigned)(ctas < (size_t)ctx->sm_count ? ctas : (size_t)ctx->sm_count);
        ProfScope prof(ctx, PROF_PARITY, stream);
        if (shape.npar == 32)
            rs255_parity_kernel<32><<<grid, 32 * kParWarps, 0, stream>>>(P);
        else if (shape.npar == 16)
            rs255_parity_kernel<16><<<grid, 32 * kParWarps, 0, stream>>>(P);
        else
            rs255_parity_kernel<8><<<grid, 32 * kParWarps, 0, stream>>>(P);
        rc = check(cudaGetLastError(), "rs255 parity launch");
    }
    if (rc != 0)
        return rc;
    RowArgs R{};
    R.parplanes = parplanes;
    R.n = n;
    R.ntiles = ntiles;
    R.parity = parity;
    R.pstride = parity_stride;
    R.bytewise = (((uintptr_t)parity | parity_stride) & 3u) != 0;
    const size_t ctas = (ntiles + kRowTeams - 1) / kRowTeams;
    const unsigned grid = (unsigned)(ctas < (size_t)ctx->sm_count ? ctas : (size_t)ctx->sm_count);
    ProfScope prof(ctx, PROF_ROWS, stream);
    if (shape.npar == 32)
        rs255_rows_kernel<32><<<grid, 32 * kRowWarps, kRowSmem, stream>>>(R);
    else if (shape.npar == 16)
        rs255_rows_kernel<16><<<grid, 32 * kRowWarps, kRowSmem, stream>>>(R);
    else
        rs255_rows_kernel<8><<<grid, 32 * kRowWarps, kRowSmem, stream>>>(R);
    return check(cudaGetLastError(), "rs255 rows launch");
}

} // namespace ppcu
