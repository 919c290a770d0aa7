// mmvq_mma.cu — the decode path for 2..8 activation columns (and small MUL_MAT_ID batches): Q4_K GEMV on the integer
// tensor-core path (mma.sync m16n8k32 u8 x s8), still HBM-bound.
//
// Why a second kernel: with dp4a every weight nibble costs one integer MAC instruction PER COLUMN, so gemv_q4k<NC> is issue-bound
// from n = 2 on (measured on the 128256x4096 output head: n = 4 2.65 TB/s, n = 8 1.24 TB/s).  Here one warp instruction
// multiplies 16 weight rows x 32 K by all 8 columns, and the per-column cost disappears.
//
// Algebra = the reference's own K-quant dot product (ggml_vec_dot_q4_K_q8_K, ggml/src/ggml-cpu/quants.c:645-718): activations
// quantized per 256 values like quantize_row_q8_K_ref (ggml/src/ggml-quants.c:2692-2729: scale 127/amax, round to nearest
// even, d8 = 1/scale), then per super-block and (row, column)
//     y += d8 * ( d * sum_j sc_j * sum_{k in j} q_k * q8_k  -  dmin * sum_j m_j * sum_{k in j} q8_k )
// with everything inside the brackets exact in int32.  sum_{k in j} q*q8 is one MMA; the sum over the mins is a second MMA
// chain whose A operand is m_j replicated over the sub-block (it accumulates over j inside the tensor core for free).
//
// Data movement:
//   * weights: every warp owns a private ring of shared-memory stages; a stage = CH super-blocks (144 B each) of 16 rows,
//     filled by ONE TMA tensor copy (a box of 16 rows x CH*144 bytes out of a tensor map over the raw quant blocks, completion
//     on an mbarrier) issued several stages ahead by one elected lane — deep prefetch without holding the data in registers
//     (first version: 16 cp.async.bulk row copies per stage; UBLKCP takes uniform registers, so the 16 lanes were
//     serialised by a software loop: 2.8 TB/s).  The MMA row <-> matrix row assignment is chosen so that the fragment loads
//     are bank-conflict-free on the dense box (row stride 144 or 288 B).
//   * activations: quantized once per CTA in the prologue (persistent CTAs) into int8 [8 columns][k + 32] (conflict-free
//     B-fragment loads) + float d8[k/256][8].
//   * K order: a dot product does not care about the order of K as long as both operands agree, so lane t of a row group
//     takes bytes 8t..8t+7 of a 32-byte group as its A fragment (one LDS.64) and the same 8 K positions of its column as B.
//   * short matrices: the super-blocks of a 16-row block are split over KS warps of the CTA (K slices) and the partial sums
//     meet in shared memory.
//   * summation order: a row's K range is ALWAYS cut into 16 leaves (super-block ranges), each leaf is summed front to back,
//     and the leaves are added along a fixed binary tree.  A warp that owns several leaves walks its subtree with a small
//     stack, the K slices of different warps are joined along the same tree — so the fp32 result of a row is bit-identical
//     for every KS, grid size and row sharding (the row-split parity property), and never depends on timing.
#include <cuda.h>
#include <cstdio>
#include <atomic>

#include "mm_problem.cuh"
#include "quantize.cuh"
#include "tc_ptx.cuh"

namespace {

constexpr int MW        = 16;                  // warps per CTA
constexpr int RB        = 16;                  // weight rows per warp task (MMA M)
constexpr int MAX_STAGE = 8;
constexpr int COL_PAD   = 32;                  // int8 column stride = k + 32
constexpr int OFF_BAR   = 64, OFF_PART = 1152, OFF_D = OFF_PART + 2 * MW * 128 * 4;      // shared-memory map (bytes)

// per weight type: bytes of one 256-value step of a row (the TMA box is `ch` of them wide), and how many values share one
// activation scale (256 like the reference's Q8_K for the K-quants, 32 like its Q8_0 / Q8_1 for the 32-blocks)
template <int TYPE> struct mma_geo;
// (STEP_B = bytes of a step in global memory, BOX_B = bytes of its TMA box / shared-memory row: Q6_K's 210-byte super-blocks are
//  fetched from the 16-byte boundary below them, 2 (sb mod 8) bytes early, in a 224-byte box)
template <> struct mma_geo<B200_TYPE_Q4_K> { static constexpr int STEP_B = SZ_Q4_K,     BOX_B = SZ_Q4_K,     ACT_BLK = 256; };
template <> struct mma_geo<B200_TYPE_Q4_0> { static constexpr int STEP_B = 8 * SZ_Q4_0, BOX_B = 8 * SZ_Q4_0, ACT_BLK = 32;  };    // 144
template <> struct mma_geo<B200_TYPE_Q8_0> { static constexpr int STEP_B = 8 * SZ_Q8_0, BOX_B = 8 * SZ_Q8_0, ACT_BLK = 32;  };    // 272
template <> struct mma_geo<B200_TYPE_Q6_K> { static constexpr int STEP_B = SZ_Q6_K,     BOX_B = 224,         ACT_BLK = 256; };
__host__ __device__ constexpr int mma_step_b(int type)  { return type == B200_TYPE_Q8_0 ? 8 * SZ_Q8_0 : (type == B200_TYPE_Q6_K ? 224 : SZ_Q4_K); }   // box bytes
__host__ __device__ constexpr int mma_act_blk(int type) { return (type == B200_TYPE_Q4_K || type == B200_TYPE_Q6_K) ? 256 : 32; }
__host__ __device__ constexpr size_t mma_off_q(int64_t k, int type)    { return (size_t) ((OFF_D + (k / mma_act_blk(type)) * 32 + 127) / 128 * 128); }
__host__ __device__ constexpr size_t mma_off_ring(int64_t k, int type, int ncol) { return (mma_off_q(k, type) + (size_t) ncol * (k + 32) + 127) / 128 * 128; }

__device__ __forceinline__ void mma_u8s8(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// the same with a zero accumulator input (no registers to clear)
__device__ __forceinline__ void mma_u8s8_z(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%10, %10, %10, %10};"
                 : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(0));
}
__device__ __forceinline__ void mma_s8s8_z(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%10, %10, %10, %10};"
                 : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(0));
}
__device__ __forceinline__ void mma16_s8s8_z(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.s32.s8.s8.s32 {%0, %1, %2, %3}, {%4, %5}, {%6}, {%7, %7, %7, %7};"
                 : "=r"(c[0]), "=r"(c[1]), "=r"(c[2]), "=r"(c[3]) : "r"(a0), "r"(a1), "r"(b0), "r"(0));
}
// sign-extended byte of x: SEL0 selects byte b (selector b + (8 | b) * 0x1110), `next` = 1 takes byte b + 1 instead
template <int SEL0> __device__ __forceinline__ int sext_byte(const uint32_t x, const int next) {
    int r;
    if (next) asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(0), "n"(SEL0 + 0x1111));
    else      asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(0), "n"(SEL0));
    return r;
}
__device__ __forceinline__ uint32_t lds_u1(const uint32_t a) {
    uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v;
}
__device__ __forceinline__ float lds_h(const uint32_t a) {           // fp16 scale of a 32-block
    uint16_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a)); return __half2float(__ushort_as_half(v));
}
// 4 / 8 bytes from a shared-memory address that is 2 mod 4 (fields of the 18- and 34-byte blocks)
__device__ __forceinline__ uint32_t lds4_odd(const uint32_t a) {
    return __byte_perm(lds_u1(a - 2), lds_u1(a + 2), 0x5432);
}
__device__ __forceinline__ uint2 lds8_odd(const uint32_t a) {
    const uint32_t w0 = lds_u1(a - 2), w1 = lds_u1(a + 2), w2 = lds_u1(a + 6);
    return make_uint2(__byte_perm(w0, w1, 0x5432), __byte_perm(w1, w2, 0x5432));
}
__device__ __forceinline__ uint2 lds_u2(const uint32_t a) {
    uint2 v; asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a)); return v;
}
__device__ __forceinline__ uint4 lds_u4(const uint32_t a) {
    uint4 v; asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a)); return v;
}

// 1-D bulk copy global -> shared, bytes reported to an mbarrier (both addresses and the size are multiples of 16)
__device__ __forceinline__ void bulk_g2s(const uint32_t dst, const void * src, const uint32_t bytes, const uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

struct mma_args {
    int ks;            // K slices per row block (1, 2, 4, 8 or 16)
    int ks_log2;
    int stages;        // ring depth per warp
    int ch;            // super-blocks per stage (1 or 2)
    int ncol;          // activation columns held in shared memory (1..8): long rows only fit with fewer than 8
    const uint8_t * img;   // quantized activations prepared by mma_quant_kernel (the shared-memory bytes [OFF_D, ring) verbatim), or NULL:
                           // every CTA quantizes for itself in its prologue
    int late;          // X / ids are not written by a kernel that may still run: the dependency wait moves to just before the first
                       // store, and this kernel's prologue and streaming overlap the previous kernel's tail (see mmvq.cu / abi.cu)
};

// adds leaf / subtree sum number i (0-based, in order) of a binary summation tree: on return after the last one of a
// power-of-two sequence, `cur` is the total
__device__ __forceinline__ void tree_push(float (&st)[4][4], const int i, float (&cur)[4]) {
    bool active = true;                 // (no early exit: the levels stay statically indexed, i.e. in registers)
#pragma unroll
    for (int lvl = 0; lvl < 4; ++lvl) {
        const bool bit = (i >> lvl) & 1;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            if (active && bit)  cur[e] = st[lvl][e] + cur[e];
            if (active && !bit) st[lvl][e] = cur[e];
        }
        active = active && bit;
    }
}

// walks a warp's (row block, super-block) steps in order: units u = warp, warp + 16, ... of this CTA; unit = (row block, K slice)
struct step_cursor {
    int u, sb, sb_end;
    int64_t rb;
    // (ks is a power of two and passed as its log2: integer divisions are slow enough to show up in a 5 us kernel)
    __device__ __forceinline__ void seek(const int units, const int ksl2, const int nsb, const int64_t rb_lo) {
        // first non-empty unit at or after u
        for (; u < units; u += MW) {
            const int q = u >> ksl2, s = u - (q << ksl2);
            sb = (nsb * s) >> ksl2; sb_end = (nsb * (s + 1)) >> ksl2; rb = rb_lo + q;
            if (sb < sb_end) return;
        }
    }
    __device__ __forceinline__ bool done(const int units) const { return u >= units; }
    __device__ __forceinline__ void advance(const int ch, const int units, const int ksl2, const int nsb, const int64_t rb_lo) {
        sb += ch;
        if (sb >= sb_end) { u += MW; seek(units, ksl2, nsb, rb_lo); }
    }
};

// ---- one 256-value step of 16 rows: st_a = the step's bytes of row 0 in the stage, r0off / r1off = offsets of this lane's two
// rows, bq_sb = this lane's B bytes of the step (column g, + 8 t), sd_sb = activation scales of columns 2t, 2t+1 for the step ----
template <int TYPE> struct mma_step;

template <> struct mma_step<B200_TYPE_Q4_K> {
    static __device__ __forceinline__ void run(const uint32_t st_a, const uint32_t r0off, const uint32_t r1off, const int t,
                                               const uint32_t bq_sb, const uint32_t sd_sb, float (&pf)[4]) {
                        const uint4 h0 = lds_u4(st_a + r0off), h1 = lds_u4(st_a + r1off);
                        // 6-bit scales / mins of the 8 sub-blocks as bytes (get_scale_min_k4, ggml/src/ggml-quants.c:818-825)
                        const uint32_t sc0[2] = {h0.y & 0x3f3f3f3fu, (h0.w & 0x0f0f0f0fu) | ((h0.y >> 2) & 0x30303030u)};
                        const uint32_t mn0[2] = {h0.z & 0x3f3f3f3fu, ((h0.w >> 4) & 0x0f0f0f0fu) | ((h0.z >> 2) & 0x30303030u)};
                        const uint32_t sc1[2] = {h1.y & 0x3f3f3f3fu, (h1.w & 0x0f0f0f0fu) | ((h1.y >> 2) & 0x30303030u)};
                        const uint32_t mn1[2] = {h1.z & 0x3f3f3f3fu, ((h1.w >> 4) & 0x0f0f0f0fu) | ((h1.z >> 2) & 0x30303030u)};
                        int al[4] = {0, 0, 0, 0}, ah[4] = {0, 0, 0, 0}, cm[4] = {0, 0, 0, 0};
                        const uint32_t bsb = bq_sb;
        #pragma unroll
                        for (int jj = 0; jj < 4; ++jj) {
                            const uint2 w0 = lds_u2(st_a + r0off + 16 + 32 * jj + 8 * t), w1 = lds_u2(st_a + r1off + 16 + 32 * jj + 8 * t);
        #pragma unroll
                            for (int hi = 0; hi < 2; ++hi) {
                                const int j = 2 * jj + hi;
                                const uint2 b = lds_u2(bsb + 32 * j);
                                const uint32_t mask = hi ? 0xF0F0F0F0u : 0x0F0F0F0Fu;      // high nibbles stay in place: 16 q, undone below
                                int c[4];
                                mma_u8s8_z(c, w0.x & mask, w1.x & mask, w0.y & mask, w1.y & mask, b.x, b.y);
                                const int s0 = (int) __byte_perm(sc0[j >> 2], 0u, 0x4440u + (j & 3));
                                const int s1 = (int) __byte_perm(sc1[j >> 2], 0u, 0x4440u + (j & 3));
                                if (hi) { ah[0] += c[0] * s0; ah[1] += c[1] * s0; ah[2] += c[2] * s1; ah[3] += c[3] * s1; }
                                else    { al[0] += c[0] * s0; al[1] += c[1] * s0; al[2] += c[2] * s1; al[3] += c[3] * s1; }
                                const uint32_t m0 = __byte_perm(mn0[j >> 2], 0u, 0x1111u * (j & 3));
                                const uint32_t m1 = __byte_perm(mn1[j >> 2], 0u, 0x1111u * (j & 3));
                                mma_u8s8(cm, m0, m1, m0, m1, b.x, b.y);
                            }
                        }
                        const uint2  d8u = lds_u2(sd_sb);
                        const float  d8a = __uint_as_float(d8u.x), d8b = __uint_as_float(d8u.y);
                        const float2 dm0 = __half22float2(*reinterpret_cast<const __half2 *>(&h0.x));
                        const float2 dm1 = __half22float2(*reinterpret_cast<const __half2 *>(&h1.x));
                        pf[0] += d8a * (dm0.x * (float) (al[0] + (ah[0] >> 4)) - dm0.y * (float) cm[0]);
                        pf[1] += d8b * (dm0.x * (float) (al[1] + (ah[1] >> 4)) - dm0.y * (float) cm[1]);
                        pf[2] += d8a * (dm1.x * (float) (al[2] + (ah[2] >> 4)) - dm1.y * (float) cm[2]);
                        pf[3] += d8b * (dm1.x * (float) (al[3] + (ah[3] >> 4)) - dm1.y * (float) cm[3]);
    }
};

// Q8_0 (block_q8_0, ggml-common.h:241-246): 8 blocks of {fp16 d; int8 qs[32]} per step; block j of a row sits at 34 j (its qs at
// 34 j + 2: 4-byte aligned for odd j only).  Per block: one s8 x s8 MMA, then y += float(sum) * d_w * d_x like
// ggml_vec_dot_q8_0_q8_0 (ggml-cpu/quants.c:400-428) with the same per-32 activation scale.
template <> struct mma_step<B200_TYPE_Q8_0> {
    static __device__ __forceinline__ void run(const uint32_t st_a, const uint32_t r0off, const uint32_t r1off, const int t,
                                               const uint32_t bq_sb, const uint32_t sd_sb, float (&pf)[4]) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t a0 = st_a + r0off + 34 * j, a1 = st_a + r1off + 34 * j;
            uint2 w0, w1;
            if (j & 1) { w0 = make_uint2(lds_u1(a0 + 2 + 8 * t), lds_u1(a0 + 6 + 8 * t)); w1 = make_uint2(lds_u1(a1 + 2 + 8 * t), lds_u1(a1 + 6 + 8 * t)); }
            else       { w0 = lds8_odd(a0 + 2 + 8 * t); w1 = lds8_odd(a1 + 2 + 8 * t); }
            const uint2 b = lds_u2(bq_sb + 32 * j);
            int c[4];
            mma_s8s8_z(c, w0.x, w1.x, w0.y, w1.y, b.x, b.y);
            const float dw0 = lds_h(a0), dw1 = lds_h(a1);
            const uint2 du = lds_u2(sd_sb + 32 * j);
            const float da = __uint_as_float(du.x), db = __uint_as_float(du.y);
            pf[0] += (float) c[0] * (dw0 * da); pf[1] += (float) c[1] * (dw0 * db);
            pf[2] += (float) c[2] * (dw1 * da); pf[3] += (float) c[3] * (dw1 * db);
        }
    }
};

// Q4_0 (block_q4_0, ggml-common.h:184-189): 8 blocks of {fp16 d; u8 qs[16]} per step; byte i of a block holds value i (low nibble)
// and value i + 16 (high nibble), each minus 8.  Lane t takes bytes 4t..4t+3: values 4t..4t+3 and 16+4t..16+4t+3 — the
// activations are stored in that order, so B is again one LDS.64.
template <> struct mma_step<B200_TYPE_Q4_0> {
    static __device__ __forceinline__ void run(const uint32_t st_a, const uint32_t r0off, const uint32_t r1off, const int t,
                                               const uint32_t bq_sb, const uint32_t sd_sb, float (&pf)[4]) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t a0 = st_a + r0off + 18 * j, a1 = st_a + r1off + 18 * j;
            const uint32_t w0 = (j & 1) ? lds_u1(a0 + 2 + 4 * t) : lds4_odd(a0 + 2 + 4 * t);
            const uint32_t w1 = (j & 1) ? lds_u1(a1 + 2 + 4 * t) : lds4_odd(a1 + 2 + 4 * t);
            const uint2 b = lds_u2(bq_sb + 32 * j);
            int c[4];
            mma_s8s8_z(c, __vsub4(w0 & 0x0F0F0F0Fu, 0x08080808u), __vsub4(w1 & 0x0F0F0F0Fu, 0x08080808u),
                          __vsub4((w0 >> 4) & 0x0F0F0F0Fu, 0x08080808u), __vsub4((w1 >> 4) & 0x0F0F0F0Fu, 0x08080808u), b.x, b.y);
            const float dw0 = lds_h(a0), dw1 = lds_h(a1);
            const uint2 du = lds_u2(sd_sb + 32 * j);
            const float da = __uint_as_float(du.x), db = __uint_as_float(du.y);
            pf[0] += (float) c[0] * (dw0 * da); pf[1] += (float) c[1] * (dw0 * db);
            pf[2] += (float) c[2] * (dw1 * da); pf[3] += (float) c[3] * (dw1 * db);
        }
    }
};

// Q6_K (block_q6_K, ggml-common.h:352-358; layout dequantize_row_q6_K, ggml-quants.c:1877-1906): value l of 32-group r of half n is
// (nibble r>>1 of ql[64n + 32(r&1) + l]) | (bits 2r, 2r+1 of qh[32n + l]) << 4, minus 32, times the int8 scale sc[8n + 2r + l/16].
// One s8 x s8 MMA (K = 16) per scale group; sum_j sc_j * (group sum) stays in int32 like ggml_vec_dot_q6_K_q8_K
// (ggml-cpu/quants.c:800-853).  Lane t takes bytes 4t..4t+3 of every 16-byte run; the activations of a 32-group are stored in the
// word order (0, 4, 1, 5, 2, 6, 3, 7), so B for the two scale groups of (n, r) is one LDS.64.  The super-block is only 2-byte
// aligned in its shared-memory row: ODD = its address is 2 mod 4 (every field offset used here is a multiple of 4).
template <> struct mma_step<B200_TYPE_Q6_K> {
    template <bool ODD> static __device__ __forceinline__ uint32_t ld(const uint32_t a) { return ODD ? lds4_odd(a) : lds_u1(a); }
    template <int R> static __device__ __forceinline__ uint32_t q6(const uint32_t la, const uint32_t lb, const uint32_t h) {
        const uint32_t lo = R == 0 ? la : (R == 1 ? lb : (R == 2 ? la >> 4 : lb >> 4));
        const uint32_t hi = R == 0 ? h << 4 : (R == 1 ? h << 2 : (R == 2 ? h : h >> 2));
        return __vsub4((lo & 0x0F0F0F0Fu) | (hi & 0x30303030u), 0x20202020u);
    }
    template <bool ODD, int R>
    static __device__ __forceinline__ void group(const uint32_t (&la0)[2], const uint32_t (&lb0)[2], const uint32_t (&h0)[2], const uint32_t (&la1)[2],
                                                 const uint32_t (&lb1)[2], const uint32_t (&h1)[2], const uint32_t (&scw0)[4], const uint32_t (&scw1)[4],
                                                 const int n, const uint32_t bq_sb, int (&acc)[4]) {
        const uint2 b = lds_u2(bq_sb + 128 * n + 32 * R);
#pragma unroll
        for (int is = 0; is < 2; ++is) {
            const int j = 8 * n + 2 * R + is;
            int c[4];
            mma16_s8s8_z(c, q6<R>(la0[is], lb0[is], h0[is]), q6<R>(la1[is], lb1[is], h1[is]), is ? b.y : b.x);
            // sign-extended byte j & 3 (prmt with the sign-replicate bit; the __byte_perm intrinsic drops that bit)
            const int s0 = sext_byte<0x8880 + 0x1110 * (R & 1) * 2 + (R & 1) * 2>(scw0[j >> 2], is);
            const int s1 = sext_byte<0x8880 + 0x1110 * (R & 1) * 2 + (R & 1) * 2>(scw1[j >> 2], is);
            acc[0] += c[0] * s0; acc[1] += c[1] * s0; acc[2] += c[2] * s1; acc[3] += c[3] * s1;
        }
    }
    template <bool ODD>
    static __device__ __forceinline__ void run_t(const uint32_t R0, const uint32_t R1, const int t, const uint32_t bq_sb, const uint32_t sd_sb, float (&pf)[4]) {
        uint32_t scw0[4], scw1[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { scw0[i] = ld<ODD>(R0 + 192 + 4 * i); scw1[i] = ld<ODD>(R1 + 192 + 4 * i); }
        int acc[4] = {0, 0, 0, 0};
#pragma unroll
        for (int n = 0; n < 2; ++n) {
            uint32_t la0[2], lb0[2], h0[2], la1[2], lb1[2], h1[2];
#pragma unroll
            for (int is = 0; is < 2; ++is) {
                la0[is] = ld<ODD>(R0 + 64 * n + 16 * is + 4 * t);      la1[is] = ld<ODD>(R1 + 64 * n + 16 * is + 4 * t);
                lb0[is] = ld<ODD>(R0 + 64 * n + 32 + 16 * is + 4 * t); lb1[is] = ld<ODD>(R1 + 64 * n + 32 + 16 * is + 4 * t);
                h0[is]  = ld<ODD>(R0 + 128 + 32 * n + 16 * is + 4 * t); h1[is] = ld<ODD>(R1 + 128 + 32 * n + 16 * is + 4 * t);
            }
            group<ODD, 0>(la0, lb0, h0, la1, lb1, h1, scw0, scw1, n, bq_sb, acc);
            group<ODD, 1>(la0, lb0, h0, la1, lb1, h1, scw0, scw1, n, bq_sb, acc);
            group<ODD, 2>(la0, lb0, h0, la1, lb1, h1, scw0, scw1, n, bq_sb, acc);
            group<ODD, 3>(la0, lb0, h0, la1, lb1, h1, scw0, scw1, n, bq_sb, acc);
        }
        const float d0 = lds_h(R0 + 208), d1 = lds_h(R1 + 208);
        const uint2 du = lds_u2(sd_sb);
        const float da = __uint_as_float(du.x), db = __uint_as_float(du.y);
        pf[0] += da * (d0 * (float) acc[0]); pf[1] += db * (d0 * (float) acc[1]);
        pf[2] += da * (d1 * (float) acc[2]); pf[3] += db * (d1 * (float) acc[3]);
    }
    static __device__ __forceinline__ void run(const uint32_t st_a, const uint32_t r0off, const uint32_t r1off, const int t,
                                               const uint32_t bq_sb, const uint32_t sd_sb, float (&pf)[4]) {
        if (st_a & 2u) run_t<true>(st_a + r0off, st_a + r1off, t, bq_sb, sd_sb, pf);       // warp-uniform
        else           run_t<false>(st_a + r0off, st_a + r1off, t, bq_sb, sd_sb, pf);
    }
};

// ---- activation quantization: the columns x[0..nc) -> int8 image q [ncol][k + 32] + scales d [k / ACT_BLK][8] ----------------
// Runs either as the GEMV's prologue (every CTA for itself, q / d in shared memory) or once per call in its own small kernel
// that writes the same image to global memory, from where the GEMV CTAs fetch it with one bulk copy (mma_quant_kernel).
// wid / nw: index and number of the warps sharing the work.
template <int TYPE>
__device__ __forceinline__ void mma_quantize(const float * const (&x)[8], const int nc, const int ncol, const int k, int8_t * q, float * d,
                                             const int wid, const int nw, const int lane) {
    constexpr int ACT_BLK = mma_geo<TYPE>::ACT_BLK;
    const int nsb = k >> 8;
    if (ACT_BLK == 256) {
        // (column, super-block) tasks: a warp takes 8 at a time and requests all their data before the first reduction, so the
        // L2 round trips of a warp's tasks overlap instead of adding up (measured: +0.5 us per column before)
        for (int task0 = wid; task0 < ncol * nsb; task0 += 8 * nw) {
            float4 v0[8], v1[8];
    #pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int task = task0 + i * nw;
                const int col = task / nsb, sb = task - col * nsb;
                v0[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f); v1[i] = v0[i];
                if (task < ncol * nsb && col < nc) {
                    const float * xc = x[0];
    #pragma unroll
                    for (int c = 1; c < 8; ++c) if (col == c) xc = x[c];
                    const float4 * src = reinterpret_cast<const float4 *>(xc + sb * 256 + lane * 8);
                    v0[i] = __ldg(src); v1[i] = __ldg(src + 1);
                }
            }
    #pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int task = task0 + i * nw;
                if (task >= ncol * nsb) break;
                const int col = task / nsb, sb = task - col * nsb;
                uint2 packed = make_uint2(0u, 0u);
                float d8 = 0.0f;
                float amax = fmaxf(fmaxf(fmaxf(fabsf(v0[i].x), fabsf(v0[i].y)), fmaxf(fabsf(v0[i].z), fabsf(v0[i].w))),
                                   fmaxf(fmaxf(fabsf(v1[i].x), fabsf(v1[i].y)), fmaxf(fabsf(v1[i].z), fabsf(v1[i].w))));
    #pragma unroll
                for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
                if (amax > 0.0f) {
                    const float is = __fdiv_rn(127.0f, amax);
                    d8 = __fdiv_rn(1.0f, is);
                    const int q0 = __float2int_rn(__fmul_rn(is, v0[i].x)), q1 = __float2int_rn(__fmul_rn(is, v0[i].y));
                    const int q2 = __float2int_rn(__fmul_rn(is, v0[i].z)), q3 = __float2int_rn(__fmul_rn(is, v0[i].w));
                    const int q4 = __float2int_rn(__fmul_rn(is, v1[i].x)), q5 = __float2int_rn(__fmul_rn(is, v1[i].y));
                    const int q6 = __float2int_rn(__fmul_rn(is, v1[i].z)), q7 = __float2int_rn(__fmul_rn(is, v1[i].w));
                    packed.x = (uint32_t) (q0 & 0xff) | ((uint32_t) (q1 & 0xff) << 8) | ((uint32_t) (q2 & 0xff) << 16) | ((uint32_t) (q3 & 0xff) << 24);
                    packed.y = (uint32_t) (q4 & 0xff) | ((uint32_t) (q5 & 0xff) << 8) | ((uint32_t) (q6 & 0xff) << 16) | ((uint32_t) (q7 & 0xff) << 24);
                }
                if (TYPE == B200_TYPE_Q6_K) {
                // words (2 lane, 2 lane + 1) of the super-block; inside their 32-group the word order is (0, 4, 1, 5, 2, 6, 3, 7)
                const int w0 = 2 * (lane & 3), p0 = w0 < 4 ? 2 * w0 : 2 * (w0 - 4) + 1;       // w0 + 1 lands at p0 + 2
                uint32_t * blk = reinterpret_cast<uint32_t *>(q + (size_t) col * (k + COL_PAD) + sb * 256 + (lane >> 2) * 32);
                blk[p0] = packed.x; blk[p0 + 2] = packed.y;
            } else {
                *reinterpret_cast<uint2 *>(q + (size_t) col * (k + COL_PAD) + sb * 256 + lane * 8) = packed;
            }
                if (lane == 0) d[sb * 8 + col] = d8;
            }
        }
    } else {
        // per-32 scales, the values of quantize_row_q8_1_ref (ggml-quants.c:260-293): 8 lanes per block, 8 blocks per lane in
        // flight.  Q4_0 stores the 8 words of a block in the order (0, 4, 1, 5, 2, 6, 3, 7) — what its nibble layout pairs up.
        const int nbk = k >> 5, total = nc * nbk, sub = lane & 7;
        for (int i = wid * 32 + lane; i < (8 - nc) * nbk; i += nw * 32) {          // dead columns: scale 0 (their int8 values are never used)
            const int col = nc + i / nbk, ib = i - (i / nbk) * nbk;
            d[ib * 8 + col] = 0.0f;
        }
        for (int blk0 = wid * 4 + (lane >> 3); blk0 < total; blk0 += 8 * nw * 4) {     // warp-uniform trip count: total is a multiple of 8
            float4 v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int blk = blk0 + i * nw * 4, bb = blk < total ? blk : total - 1;
                const int col = bb / nbk, ib = bb - col * nbk;
                const float * xc = x[0];
#pragma unroll
                for (int c = 1; c < 8; ++c) if (col == c) xc = x[c];
                v[i] = __ldg(reinterpret_cast<const float4 *>(xc + ib * 32 + sub * 4));
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int blk = blk0 + i * nw * 4, bb = blk < total ? blk : total - 1;
                const int col = bb / nbk, ib = bb - col * nbk;
                const q8_part qp = quantize_q8_1_part(v[i]);
                if (blk < total) {
                    const int w = TYPE == B200_TYPE_Q4_0 ? (sub < 4 ? 2 * sub : 2 * (sub - 4) + 1) : sub;
                    *reinterpret_cast<uint32_t *>(q + (size_t) col * (k + COL_PAD) + ib * 32 + 4 * w) = qp.packed;
                    if (sub == 0) d[ib * 8 + col] = qp.d;
                }
            }
        }
    }
}

// the activations of one MUL_MAT call (<= 8 columns, no batch dims), quantized ONCE for all CTAs of the GEMV that follows
template <int TYPE>
__global__ void __launch_bounds__(256) mma_quant_kernel(const mm_problem P, const int ncol, uint8_t * img) {
    // X may be produced by the previous kernel of the stream, and the previous GEMV may still be fetching the image this kernel
    // rewrites: wait for it; the GEMV that consumes the image may be launched meanwhile (it waits for this grid before it reads)
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int k = (int) P.k, nc = (int) P.ne11;
    const float * xs[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) xs[c] = reinterpret_cast<const float *>(reinterpret_cast<const char *>(P.X) + (c < nc ? c : 0) * P.nb11);
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    mma_quantize<TYPE>(xs, nc, ncol, k, reinterpret_cast<int8_t *>(img + (mma_off_q(k, TYPE) - OFF_D)), reinterpret_cast<float *>(img),
                       gtid >> 5, (int) (gridDim.x * blockDim.x) >> 5, threadIdx.x & 31);
}

#ifdef MMA_TRACE_ENABLE      // development aid: phase time stamps of CTA 0 / warp 0 (build with -DMMA_TRACE_ENABLE)
#define MMA_TRACE(i) do { if (blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); tr_[i] = t_; } } while (0)
#else
#define MMA_TRACE(i) do { } while (0)
#endif

template <int TYPE>
__global__ void __launch_bounds__(MW * 32, 1)
gemv_q4k_mma(const __grid_constant__ CUtensorMap tmW, const mm_problem P, const mma_args A) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int k = (int) P.k, nsb = k >> 8;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
#ifdef MMA_TRACE_ENABLE
    unsigned long long tr_[8] = {};
#endif
    MMA_TRACE(0);
    // shared memory map
    float **  s_y    = reinterpret_cast<float **>(smem);                              // 8 output column pointers
    float *   s_part = reinterpret_cast<float *>(smem + OFF_PART);                    // [2][MW][128] K-slice partial sums
    constexpr int STEP_B = mma_geo<TYPE>::STEP_B, BOX_B = mma_geo<TYPE>::BOX_B, ACT_BLK = mma_geo<TYPE>::ACT_BLK;
    float *   s_d    = reinterpret_cast<float *>(smem + OFF_D);                       // [k / ACT_BLK][8] activation scales
    int8_t *  s_q    = reinterpret_cast<int8_t *>(smem + mma_off_q(k, TYPE));         // [8][k + 32]
    const uint32_t stage_b = (uint32_t) A.ch * RB * BOX_B;                            // [MW][stages][16 rows][ch * BOX_B]
    const uint32_t ring0 = tc::smem_u32(smem) + (uint32_t) mma_off_ring(k, TYPE, A.ncol) + (uint32_t) (warp * A.stages) * stage_b;
    const uint32_t bar0  = tc::smem_u32(smem) + OFF_BAR + (uint32_t) warp * MAX_STAGE * 8;

    const uint32_t barI = tc::smem_u32(smem) + OFF_BAR + MW * MAX_STAGE * 8;     // completion of the activation image copy
    if (tid == 0) { tc::tma_prefetch_desc(&tmW); tc::mbar_init(barI, 1); }
    if (lane == 0) {
        for (int s = 0; s < A.stages; ++s) tc::mbar_init(bar0 + 8 * s, 1);
        tc::mbar_fence_init();
    }
    __syncwarp();

    // this CTA's share of the row blocks (persistent: the activations are quantized once per CTA)
    const uint32_t nrb32 = (uint32_t) ((P.m + RB - 1) >> 4);                       // m <= 2^30: 32-bit arithmetic, no 64-bit divisions
    const int64_t  rb_lo = nrb32 < (1u << 20) ? (int64_t) (nrb32 * blockIdx.x / gridDim.x) : (int64_t) nrb32 * blockIdx.x / gridDim.x;
    const int64_t  rb_hi = nrb32 < (1u << 20) ? (int64_t) (nrb32 * (blockIdx.x + 1) / gridDim.x) : (int64_t) nrb32 * (blockIdx.x + 1) / gridDim.x;
    const int     units = (int) (rb_hi - rb_lo) * A.ks;
    const int     rounds = (units + MW - 1) / MW;

    mm_cols<8> C;
    bool live = true;
    // one ring fill: super-blocks [sb, sb + ch) of rows [16 rb, 16 rb + 16) of matrix `mat` (rows / K past the end are zero-filled)
    int mat = 0;
    auto fill = [&](const step_cursor & c, const int stage) {
        const uint32_t bar = bar0 + 8 * stage;
        if (tc::elect_one()) {
            tc::mbar_arrive_expect_tx(bar, stage_b);
            // (Q6_K: the box starts at the 16-byte boundary below the super-block, 2 (sb mod 8) bytes = (sb mod 8) elements early)
            tc::tma_load_3d(ring0 + stage * stage_b, &tmW, bar, c.sb * (STEP_B / 2) - (TYPE == B200_TYPE_Q6_K ? (c.sb & 7) : 0), (int) (c.rb * RB), mat);
        }
        __syncwarp();
    };
    step_cursor pc;                       // producer cursor
    int p_stage = 0, in_flight = 0;
    auto start_fills = [&]() {
        pc.u = warp; pc.seek(units, A.ks_log2, nsb, rb_lo);
        for (; in_flight < A.stages && !pc.done(units); ++in_flight) {
            fill(pc, p_stage);
            if (++p_stage == A.stages) p_stage = 0;
            pc.advance(A.ch, units, A.ks_log2, nsb, rb_lo);
        }
    };
    auto set_mat = [&]() { mat = P.nb02 > 0 ? (int) ((C.W - P.W) / P.nb02) : 0; };

    // weights are never written by a kernel of this path: in MM_PLAIN the first ring fills go out before the dependency wait
    if (P.mode == MM_PLAIN) {
        if (gridDim.y == 1) {               // one matrix, <= 8 columns: nothing to resolve (mm_resolve costs several 64-bit divisions)
            C.W = P.W; C.nc = (int) P.ne11;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const int cc = c < C.nc ? c : 0;
                C.plane_col[c] = cc;
                C.x[c] = reinterpret_cast<const float *>(reinterpret_cast<const char *>(P.X) + cc * P.nb11);
                C.y[c] = P.Y + cc * P.m;
            }
        } else {
            live = mm_resolve<8>(P, blockIdx.y, C);
        }
        set_mat();
        start_fills();
    }
    MMA_TRACE(1);
    if (A.late) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    else        asm volatile("griddepcontrol.wait;" ::: "memory");
    MMA_TRACE(2);
    if (P.mode != MM_PLAIN) {
        live = mm_resolve<8>(P, blockIdx.y, C);
        set_mat();
        if (live) start_fills();
    }
    if (!live) return;                    // block-uniform; nothing was issued

    // ---- prologue: the (up to 8) activation columns -> int8 + one scale per 256 values -----------------------------------
    if (tid < 8) {
        float * y = C.y[0];
#pragma unroll
        for (int c = 1; c < 8; ++c) if (tid == c) y = C.y[c];
        s_y[tid] = y;
    }
    if (A.img) {
        // quantized once by mma_quant_kernel (complete: the dependency wait above was not deferred): one bulk copy per CTA
        const uint32_t bytes = (uint32_t) (mma_off_ring(k, TYPE, A.ncol) - OFF_D);
        if (tid == 0) {
            tc::mbar_arrive_expect_tx(barI, bytes);
            bulk_g2s(tc::smem_u32(smem) + OFF_D, A.img, bytes, barI);
        }
        __syncthreads();                     // barI's initialisation is visible to everybody
        tc::mbar_wait(barI, 0);
    } else {
        const float * xs[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) xs[c] = C.x[c];
        mma_quantize<TYPE>(xs, C.nc, A.ncol, k, s_q, s_d, warp, MW, lane);
    }
    MMA_TRACE(3);
    __syncthreads();
    MMA_TRACE(4);
    if (!A.late) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    bool waited = !A.late;

    // ---- main loop -------------------------------------------------------------------------------------------------------
    // (lanes of columns that are not held re-read the last one; those results are never stored)
    const uint32_t bq = tc::smem_u32(s_q) + (uint32_t) (g < A.ncol ? g : A.ncol - 1) * (uint32_t) (k + COL_PAD) + 8u * t;   // column g, K bytes 8t..8t+7 of a sub-block
    const uint32_t sd = tc::smem_u32(s_d) + 8u * t;                                            // d8 of columns 2t, 2t+1
    // MMA row group g <-> rows rho, rho + 8 of the block: on the dense box (row stride 36 or 72 words) the LDS.64 of a
    // half-warp (4 row groups x 4 lanes x 8 B) must hit 32 distinct banks: stride 72 words = 8 mod 32 works with rho = g,
    // stride 36 words = 4 mod 32 needs the rows of a half-warp two apart
    const int      rho   = A.ch == 1 ? 2 * (g & 3) + (g >> 2) : g;
    const uint32_t rstr  = (uint32_t) A.ch * BOX_B;
    const uint32_t r0off = (uint32_t) rho * rstr, r1off = (uint32_t) (rho + 8) * rstr;
    int      c_stage = 0;
    uint32_t c_phase = 0;
    for (int round = 0; round < rounds; ++round) {
        const int u = round * MW + warp;
        float accf[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        int64_t rb = 0;
        int     ksl = 0;
        if (u < units) {
            const int q = u >> A.ks_log2;
            ksl = u - (q << A.ks_log2);
            rb  = rb_lo + q;
            const int lpw = 16 >> A.ks_log2;                                  // summation-tree leaves owned by this warp
            float st[4][4] = {};
            const int sb_last = ((nsb * (ksl + 1)) >> A.ks_log2) - 1;             // last super-block of the unit
            int in_chunk = 0;
            for (int li = 0; li < lpw; ++li) {
              const int lf = ksl * lpw + li;
              float pf[4] = {0.0f, 0.0f, 0.0f, 0.0f};
              for (int sb = (nsb * lf) >> 4; sb < (nsb * (lf + 1)) >> 4; ++sb) {
                if (in_chunk == 0) tc::mbar_wait(bar0 + 8 * c_stage, c_phase);
                const uint32_t st_a = ring0 + c_stage * stage_b + in_chunk * BOX_B + (TYPE == B200_TYPE_Q6_K ? 2u * (sb & 7) : 0u);
                mma_step<TYPE>::run(st_a, r0off, r1off, t, bq + 256u * sb, sd + (uint32_t) (256 / ACT_BLK) * 32u * sb, pf);
                if (++in_chunk == A.ch || sb == sb_last) {
                    // the stage has been read (values are in registers): hand it to the next fill
                    in_chunk = 0;
                    __syncwarp();
                    if (!pc.done(units)) {
                        fill(pc, c_stage);
                        pc.advance(A.ch, units, A.ks_log2, nsb, rb_lo);
                    }
                    if (++c_stage == A.stages) { c_stage = 0; c_phase ^= 1u; }
                }
              }
              tree_push(st, li, pf);
              if (li == lpw - 1) { accf[0] = pf[0]; accf[1] = pf[1]; accf[2] = pf[2]; accf[3] = pf[3]; }
            }
        }
        if (round == 0) MMA_TRACE(5);
        // ---- end of the unit: add the K slices in slice order, store -----------------------------------------------------
        bool store = u < units;
        if (A.ks > 1) {
            float * part = s_part + (round & 1) * MW * 128;
            *reinterpret_cast<float4 *>(part + warp * 128 + lane * 4) = make_float4(accf[0], accf[1], accf[2], accf[3]);
            __syncthreads();
            store = store && ksl == 0;
            if (store) {
                float st[4][4] = {};
                for (int s = 0; s < A.ks; ++s) {               // the slices' subtree sums, joined along the same tree
                    const float4 o = *reinterpret_cast<const float4 *>(part + (warp + s) * 128 + lane * 4);
                    accf[0] = o.x; accf[1] = o.y; accf[2] = o.z; accf[3] = o.w;
                    tree_push(st, s, accf);
                }
            }
        }
        if (store) {
            if (!waited) { asm volatile("griddepcontrol.wait;" ::: "memory"); waited = true; }
            const int64_t row0 = rb * RB + rho, row1 = row0 + 8;
            const int c0 = 2 * t, c1 = c0 + 1;
            if (c0 < C.nc) {
                float * y = s_y[c0];
                if (row0 < P.m) mm_store(P, y, row0, accf[0]);
                if (row1 < P.m) mm_store(P, y, row1, accf[2]);
            }
            if (c1 < C.nc) {
                float * y = s_y[c1];
                if (row0 < P.m) mm_store(P, y, row0, accf[1]);
                if (row1 < P.m) mm_store(P, y, row1, accf[3]);
            }
        }
    }
    MMA_TRACE(6);
#ifdef MMA_TRACE_ENABLE
    if (blockIdx.x == 0 && blockIdx.y == 0 && tid == 0)
        printf("mma trace (ns from start): fills issued %llu, dependency wait done %llu, prologue done %llu, after barrier %llu, first unit done %llu, end %llu  (grid %u, ks %d, stages %d, ch %d)\n",
               tr_[1] - tr_[0], tr_[2] - tr_[0], tr_[3] - tr_[0], tr_[4] - tr_[0], tr_[5] - tr_[0], tr_[6] - tr_[0], gridDim.x, A.ks, A.stages, A.ch);
#endif
}

}  // namespace

typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static encode_tiled_fn mma_get_encode() {
    static encode_tiled_fn fn = [] {
        void * p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) p = nullptr;
        return (encode_tiled_fn) p;
    }();
    return fn;
}

// shared memory for k values per column with `stages` ring slots of `ch` steps per warp
static size_t mma_smem(int type, int64_t k, int stages, int ch, int ncol) { return mma_off_ring(k, type, ncol) + (size_t) MW * stages * ch * RB * mma_step_b(type); }

// number of weight matrices behind P.W (batches / experts), or 0 when they are not equally spaced
static int64_t mma_nmat(const mm_problem & P) {
    if (P.mode != MM_PLAIN) return P.n_expert;
    const int64_t ne02 = P.ne12 / P.r2, ne03 = P.ne13 / P.r3;
    if (ne03 > 1 && P.nb03 != P.nb02 * ne02) return 0;
    return ne02 * ne03;
}

static bool mma_type_ok(int type) { return type == B200_TYPE_Q4_K || type == B200_TYPE_Q4_0 || type == B200_TYPE_Q8_0 || type == B200_TYPE_Q6_K; }

// activation columns a tile can have: MUL_MAT n (<= 8), MUL_MAT_ID up to 8 pairs of an expert
static int mma_ncol(const mm_problem & P) { return P.mode == MM_PLAIN ? (int) (P.ne11 < 8 ? P.ne11 : 8) : 8; }

int b200_gemv_mma_ok(const b200_ctx * ctx, int type, const mm_problem & P) {
    if (!mma_type_ok(type) || P.k % 256 != 0 || P.k > (1 << 24) || P.m > (1 << 30) || !mma_get_encode()) return 0;
    if (mma_smem(type, P.k, 2, 1, mma_ncol(P)) > ctx->smem_optin) return 0;
    // TMA over the raw blocks: base and strides 16-byte aligned, the matrices of all batches / experts equally spaced
    if (((uintptr_t) P.W | (uintptr_t) P.nb01 | (uintptr_t) P.nb02 | (uintptr_t) P.nb03) & 15) return 0;
    const int64_t nmat = mma_nmat(P);
    if (nmat <= 0 || (nmat > 1 && P.nb02 < P.nb01 * P.m)) return 0;
    if (((uintptr_t) P.X | (uintptr_t) P.nb11 | (uintptr_t) P.nb12 | (uintptr_t) P.nb13) & 15) return 0;
    return 1;
}

template <int TYPE>
static int mma_quant_launch_t(b200_ctx * ctx, const mm_problem & P, int ncol, uint8_t * img) {
    const int64_t warps = mma_geo<TYPE>::ACT_BLK == 256 ? (int64_t) ncol * (P.k / 256) : (P.ne11 * (P.k / 32) + 3) / 4;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) ((warps + 7) / 8)); cfg.blockDim = dim3(256); cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = ctx->pdl ? 1 : 0;
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, mma_quant_kernel<TYPE>, P, ncol, img));
    ctx->launches++;
    return B200_OK;
}

template <int TYPE>
static int mma_launch_t(b200_ctx * ctx, const cudaLaunchConfig_t & cfg, const CUtensorMap & tmw, const mm_problem & P, const mma_args & A) {
    static std::atomic<bool> attr_done[16];   // per device (several backends of one device may launch from different threads)
    if (!attr_done[ctx->device & 15]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(gemv_q4k_mma<TYPE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) ctx->smem_optin));
        attr_done[ctx->device & 15] = true;
    }
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, gemv_q4k_mma<TYPE>, tmw, P, A));
    return B200_OK;
}

// f32 activations in P.X (quantized in the kernel prologue); up to 8 columns per tile (MM_PLAIN: ceil(ne11 / 8) tiles per batch)
// live_tiles: how many of the tiles are expected to have work (MM_ID_SORTED launches an upper bound; 0 = all)
int b200_launch_gemv_mma(b200_ctx * ctx, int type, const mm_problem & Pin, int64_t tiles, bool late_wait, int64_t live_tiles) {
    if (tiles <= 0 || Pin.m <= 0) return B200_OK;
    if (tiles > 65535) { b200_set_error("gemv_mma: %lld tiles exceed gridDim.y", (long long) tiles); return B200_E_ARG; }
    mm_problem P = Pin;
    const int64_t nmat = mma_nmat(P);
    if (nmat <= 1) P.nb02 = 0;                    // the kernel derives the matrix index from (W of the tile - P.W) / nb02
    const int nsb = (int) (P.k / 256);
    // persistent CTAs share the SMs evenly between the tiles; K slices so that every warp of every CTA has work even when the
    // matrix is short (the result does not depend on the slicing: fixed summation tree)
    const int64_t nrb = (P.m + RB - 1) / RB;
    int64_t ctas = ctx->sm_count / (live_tiles > 0 && live_tiles < tiles ? live_tiles : tiles);
    if (ctas < 1) ctas = 1;
    int ks = 1;
    while (ks < MW && 4 * nrb * ks < 3 * ctas * MW && nsb / (2 * ks) >= 1) ks *= 2;      // measured best: the first ks with ~one unit per warp
    if (ctx->mma_ks > 0) ks = ctx->mma_ks;
    if (ks != 1 && ks != 2 && ks != 4 && ks != 8 && ks != 16) { b200_set_error("gemv_mma: K slices must be 1, 2, 4, 8 or 16"); return B200_E_ARG; }
    mma_args A = {};
    A.ks = ks;
    A.ks_log2 = ks == 1 ? 0 : ks == 2 ? 1 : ks == 4 ? 2 : ks == 8 ? 3 : 4;
    A.late = ctx->pdl && late_wait ? 1 : 0;
    // steps per stage: 2 when every K slice is a whole number of pairs (the 272-byte Q8_0 step exceeds the 256-element box: 1)
    const int step_b = mma_step_b(type);
    A.ch = (nsb % (2 * ks) == 0 && step_b <= 256 && type != B200_TYPE_Q6_K) ? 2 : 1;
    if (ctx->mma_ch == 1) A.ch = 1;
    A.ncol = mma_ncol(P);
    int stages = MAX_STAGE;
    while (stages > 1 && mma_smem(type, P.k, stages, A.ch, A.ncol) > ctx->smem_optin) --stages;
    if (stages < 2 && A.ch == 2) { A.ch = 1; stages = 2; }
    if (mma_smem(type, P.k, stages, A.ch, A.ncol) > ctx->smem_optin) { b200_set_error("gemv_mma: k = %lld needs too much shared memory", (long long) P.k); return B200_E_ARG; }
    A.stages = stages;

    // raw weight blocks: [nmat][m][row bytes] seen as 16-bit elements, box = 16 rows x ch steps
    CUtensorMap tmw;
    const cuuint64_t wdim[3]    = {(cuuint64_t) P.nb01 / 2, (cuuint64_t) P.m, (cuuint64_t) (nmat > 1 ? nmat : 1)};
    const cuuint64_t wstride[2] = {(cuuint64_t) P.nb01, (cuuint64_t) (nmat > 1 ? P.nb02 : P.nb01 * P.m)};
    const cuuint32_t wbox[3]    = {(cuuint32_t) (A.ch * step_b / 2), (cuuint32_t) RB, 1};
    const cuuint32_t westr[3]   = {1, 1, 1};
    const CUresult wr = mma_get_encode()(&tmw, CU_TENSOR_MAP_DATA_TYPE_UINT16, 3, (void *) P.W, wdim, wstride, wbox, westr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (wr != CUDA_SUCCESS) { b200_set_error("gemv_mma: cuTensorMapEncodeTiled failed (%d)", (int) wr); return B200_E_CUDA; }

    // one MUL_MAT without batch dims: quantize the activations once in a small kernel instead of in every CTA's prologue (measured
    // per CTA: 3.4 us at n = 1, 7 us at n = 8 for k = 4096 — more than the streaming of a decoder-layer matrix)
    // Only when the activations are still being produced (no deferred wait possible): otherwise the CTA prologue overlaps the
    // previous kernel's tail and costs nothing (mma_prequant: 0 never, 1 when dependent, 2 always).
    // Measured on dependent chains of 4096x4096 matmuls (tools/gemv_chain.py, us per matmul, in-CTA / once): n = 8 Q8_0 13.7 / 10.7,
    // Q4_0 13.2 / 10.3, Q4_K 9.4 / 8.2, Q6_K 11.4 / 10.2; n <= 4: equal or worse (the extra kernel boundary costs what it saves).
    const bool pre_possible = ctx->mma_prequant > 0 && (ctx->mma_prequant == 2 || P.ne11 >= 5) && P.mode == MM_PLAIN && tiles == 1 && ctx->pdl;
    if (pre_possible) {
        // reserved whether or not this particular launch uses it: a later launch of the same shape may sit inside a stream capture,
        // where the scratch cannot grow
        const int rc = b200_ctx_reserve(ctx, mma_off_ring(P.k, type, A.ncol) - OFF_D);
        if (rc) return rc;
    }
    if (pre_possible && (ctx->mma_prequant == 2 || !late_wait)) {
        int rc = B200_OK;
        uint8_t * img = (uint8_t *) ctx->ws;
        switch (type) {
            case B200_TYPE_Q4_K: rc = mma_quant_launch_t<B200_TYPE_Q4_K>(ctx, P, A.ncol, img); break;
            case B200_TYPE_Q4_0: rc = mma_quant_launch_t<B200_TYPE_Q4_0>(ctx, P, A.ncol, img); break;
            case B200_TYPE_Q8_0: rc = mma_quant_launch_t<B200_TYPE_Q8_0>(ctx, P, A.ncol, img); break;
            case B200_TYPE_Q6_K: rc = mma_quant_launch_t<B200_TYPE_Q6_K>(ctx, P, A.ncol, img); break;
            default: rc = B200_E_UNSUPPORTED;
        }
        if (rc) return rc;
        A.img  = img;
        A.late = 0;                       // the image must be complete before the CTAs fetch it
    }
    const int64_t max_useful = (nrb * ks + MW - 1) / MW;
    if (ctas > max_useful) ctas = max_useful;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim          = dim3((unsigned) ctas, (unsigned) tiles);
    cfg.blockDim         = dim3(MW * 32);
    cfg.dynamicSmemBytes = mma_smem(type, P.k, stages, A.ch, A.ncol);
    cfg.stream           = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs    = attr;
    cfg.numAttrs = ctx->pdl ? 1 : 0;
    int rc = B200_E_UNSUPPORTED;
    switch (type) {
        case B200_TYPE_Q4_K: rc = mma_launch_t<B200_TYPE_Q4_K>(ctx, cfg, tmw, P, A); break;
        case B200_TYPE_Q4_0: rc = mma_launch_t<B200_TYPE_Q4_0>(ctx, cfg, tmw, P, A); break;
        case B200_TYPE_Q8_0: rc = mma_launch_t<B200_TYPE_Q8_0>(ctx, cfg, tmw, P, A); break;
        case B200_TYPE_Q6_K: rc = mma_launch_t<B200_TYPE_Q6_K>(ctx, cfg, tmw, P, A); break;
        default: b200_set_error("gemv_mma: unsupported weight type %d", type);
    }
    if (rc == B200_OK) ctx->launches++;
    return rc;
}
