// mmvq.cu — the decode path: memory-bound GEMV family for 1..8 activation columns.
//
// Replaces the hot loop of ggml_compute_forward_mul_mat[_id] (/root/reference/ggml/src/ggml-cpu/ggml-cpu.c:1151-1239,
// 1443-1511) — vec_dot(W row, quantized X column) for every (row, column) — with integer dot products against
// Q8_1-quantized activations (same d/qs as quantize_row_q8_1_ref).  The per-block algebra is
//   Q4_0: d * (d8 * sum((nib)*q8) - 8 * s8)                      (block_q4_0, ggml-common.h:184-189)
//   Q8_0: d * d8 * sum(q*q8)                                       (block_q8_0, :241-246)
//   Q4_K: d * sc_j * d8 * sum(nib*q8) - dmin * m_j * d8 * sum(q8)  (block_q4_K, :317-328; scales ggml-quants.c:818-825)
//   Q6_K: d * d8 * (sc_2j * sum_lo((q-32)*q8) + sc_2j+1 * sum_hi)  (block_q6_K, :352-358; layout ggml-quants.c:1877-1906)
//
// Two kernels:
//   gemv_generic  — any of the four types, any strides / batch / MUL_MAT_ID mode, activations pre-quantized into
//                   global planes.  One warp per weight row, one 32-weight unit per lane per step.
//   gemv_q4k      — the Q4_K fast path: persistent CTAs (one per SM), activations quantized IN the prologue into
//                   swizzled shared memory (no separate launch, no q8 round trip through HBM), weights streamed with
//                   coalesced 128-bit no-allocate loads (4 lanes per 144-byte superblock), d/dmin/6-bit scales applied
//                   in registers, one fragment of software prefetch, warp-shuffle reduction.
#include <atomic>

#include "mm_problem.cuh"
#include "quantize.cuh"
#include "ops.cuh"

// ------------------------------------------------------------------------------------------------------------------
// generic kernel
// ------------------------------------------------------------------------------------------------------------------
namespace {

struct act_unit {      // one 32-value activation block of one column
    int   a[8];
    float d, s;
};

__device__ __forceinline__ act_unit load_act(const q8_planes & A, int64_t col, int64_t ib) {
    act_unit u;
    const uint4 * q = reinterpret_cast<const uint4 *>(A.qs + col * A.k + ib * 32);
    const uint4 q0 = __ldg(q), q1 = __ldg(q + 1);
    u.a[0] = q0.x; u.a[1] = q0.y; u.a[2] = q0.z; u.a[3] = q0.w;
    u.a[4] = q1.x; u.a[5] = q1.y; u.a[6] = q1.z; u.a[7] = q1.w;
    const float2 ds = __half22float2(__ldg(A.ds + col * (A.k >> 5) + ib));
    u.d = ds.x; u.s = ds.y;
    return u;
}

// N consecutive 32-bit words from a 2-byte aligned address: N + 1 aligned word loads + funnel shifts (the block formats are
// only 2-byte aligned; two 16-bit loads per word cost twice the load instructions — the generic kernel is issue-bound).
// Reads up to 2 bytes past the last word when the address is 4-byte aligned (tensor rows are padded, ggml-cuda.cu:802-816 contract).
template <int N>
__device__ __forceinline__ void ldw_b2(const uint8_t * p, uint32_t (&out)[N]) {
    const uint32_t * a  = reinterpret_cast<const uint32_t *>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t) 3);
    const uint32_t   sh = ((uint32_t) reinterpret_cast<uintptr_t>(p) & 2u) * 8u;
    uint32_t w[N + 1];
#pragma unroll
    for (int i = 0; i <= N; ++i) w[i] = __ldg(a + i);
#pragma unroll
    for (int i = 0; i < N; ++i) out[i] = __funnelshift_r(w[i], w[i + 1], sh);
}

// weights of one 32-element unit, unpacked to int8x4 words + the floats needed to scale the integer sums
template <int TYPE> struct w_unit;

template <> struct w_unit<B200_TYPE_Q4_0> {
    int lo[4], hi[4]; float d;
    __device__ __forceinline__ void load(const uint8_t * row, int64_t ib) {
        const uint8_t * bp = row + ib * SZ_Q4_0;
        d = ld_half(bp);
        uint32_t q[4];
        ldw_b2<4>(bp + 2, q);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            lo[i] = q[i] & 0x0F0F0F0Fu;
            hi[i] = (q[i] >> 4) & 0x0F0F0F0Fu;
        }
    }
    __device__ __forceinline__ float dot(const act_unit & u) const {
        int sumi = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) { sumi = __dp4a(lo[i], u.a[i], sumi); sumi = __dp4a(hi[i], u.a[i + 4], sumi); }
        return d * (u.d * (float) sumi - 8.0f * u.s);
    }
};

template <> struct w_unit<B200_TYPE_Q8_0> {
    int v[8]; float d;
    __device__ __forceinline__ void load(const uint8_t * row, int64_t ib) {
        const uint8_t * bp = row + ib * SZ_Q8_0;
        d = ld_half(bp);
        uint32_t q[8];
        ldw_b2<8>(bp + 2, q);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = (int) q[i];
    }
    __device__ __forceinline__ float dot(const act_unit & u) const {
        int sumi = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) sumi = __dp4a(v[i], u.a[i], sumi);
        return d * u.d * (float) sumi;
    }
};

template <> struct w_unit<B200_TYPE_Q4_K> {
    int v[8]; float dsc, dmn;
    __device__ __forceinline__ void load(const uint8_t * row, int64_t ib) {
        const int64_t sb = ib >> 3;
        const int     j  = (int) (ib & 7);
        const uint32_t * bp = reinterpret_cast<const uint32_t *>(row + sb * SZ_Q4_K);     // 4-byte aligned
        const float2 dm = __half22float2(*reinterpret_cast<const half2 *>(bp));
        int sc, mn;
        q4k_scale_min(j, __ldg(bp + 1), __ldg(bp + 2), __ldg(bp + 3), sc, mn);
        dsc = dm.x * (float) sc;
        dmn = dm.y * (float) mn;
        const uint32_t * q = bp + 4 + 8 * (j >> 1);
        const int sh = 4 * (j & 1);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = (__ldg(q + i) >> sh) & 0x0F0F0F0Fu;
    }
    __device__ __forceinline__ float dot(const act_unit & u) const {
        int sumi = 0, sum8 = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) { sumi = __dp4a(v[i], u.a[i], sumi); sum8 = __dp4a(0x01010101, u.a[i], sum8); }
        return u.d * (dsc * (float) sumi - dmn * (float) sum8);
    }
};

template <> struct w_unit<B200_TYPE_Q6_K> {
    int v[8]; float d0, d1;
    __device__ __forceinline__ void load(const uint8_t * row, int64_t ib) {
        const int64_t sb = ib >> 3;
        const int     j  = (int) (ib & 7);
        const int     nh = j >> 2, r = j & 3;
        const uint8_t * bp = row + sb * SZ_Q6_K;                                            // 2-byte aligned only
        const uint8_t * ql = bp + nh * 64 + (r & 1) * 32;
        const uint8_t * qh = bp + 128 + nh * 32;
        const uint16_t scs = __ldg(reinterpret_cast<const uint16_t *>(bp + 192 + 2 * j));
        const float d = ld_half(bp + 208);
        d0 = d * (float) (int8_t) (scs & 0xff);
        d1 = d * (float) (int8_t) (scs >> 8);
        const int shl = 4 * (r >> 1), shh = 2 * r;
        uint32_t lw[8], hw[8];
        ldw_b2<8>(ql, lw);
        ldw_b2<8>(qh, hw);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t l = (lw[i] >> shl) & 0x0F0F0F0Fu;
            const uint32_t h = ((hw[i] >> shh) & 0x03030303u) << 4;
            v[i] = __vsubss4(l | h, 0x20202020);
        }
    }
    __device__ __forceinline__ float dot(const act_unit & u) const {
        int s0 = 0, s1 = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) { s0 = __dp4a(v[i], u.a[i], s0); s1 = __dp4a(v[i + 4], u.a[i + 4], s1); }
        return u.d * (d0 * (float) s0 + d1 * (float) s1);
    }
};

// N4 — Q4_1 / Q5_0 / Q5_1 / Q5_K (ggml-common.h:191-239, 330-346; dequantize_row_* ggml-quants.c:417-490, 1669-1697): the same per-block algebra
// with a block minimum (x = d q + m: the minimum meets the activations' block sum s) and / or a fifth bit (bit e of qh belongs to element e;
// Q5_K: bit j of qh[l] belongs to sub-block j).
__device__ __forceinline__ uint32_t spread4(const uint32_t h4) {          // bits 0..3 -> bit 4 of bytes 0..3
    return ((h4 & 1u) << 4) | ((h4 & 2u) << 11) | ((h4 & 4u) << 18) | ((h4 & 8u) << 25);
}

template <> struct w_unit<B200_TYPE_Q4_1> {
    int lo[4], hi[4]; float d, m;
    __device__ __forceinline__ void load(const uint8_t * row, int64_t ib) {
        const uint8_t * bp = row + ib * SZ_Q4_1;
        d = ld_half(bp); m = ld_half(bp + 2);
        uint32_t q[4];
        ldw_b2<4>(bp + 4, q);
#pragma unroll
        for (int i = 0; i < 4; ++i) { lo[i] = q[i] & 0x0F0F0F0Fu; hi[i] = (q[i] >> 4) & 0x0F0F0F0Fu; }
    }
    __device__ __forceinline__ float dot(const act_unit & u) const {
        int sumi = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) { sumi = __dp4a(lo[i], u.a[i], sumi); sumi = __dp4a(hi[i], u.a[i + 4], sumi); }
        return d * u.d * (float) sumi + m * u.s;
    }
};

template <int TYPE> struct w_unit_q5 {          // Q5_0 (TYPE = Q5_0: offset 16, no minimum) and Q5_1 (minimum)
    int v[8]; float d, m;
    __device__ __forceinline__ void load(const uint8_t * row, int64_t ib) {
        constexpr int SZ = TYPE == B200_TYPE_Q5_0 ? SZ_Q5_0 : SZ_Q5_1, OFF = TYPE == B200_TYPE_Q5_0 ? 2 : 4;
        const uint8_t * bp = row + ib * SZ;
        d = ld_half(bp); m = TYPE == B200_TYPE_Q5_0 ? 0.0f : ld_half(bp + 2);
        uint32_t w[5];
        ldw_b2<5>(bp + OFF, w);                  // qh, then the 16 nibble bytes
        const uint32_t qh = w[0];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            v[i]     = (int) ((w[i + 1] & 0x0F0F0F0Fu)        | spread4((qh >> (4 * i)) & 0xFu));
            v[i + 4] = (int) (((w[i + 1] >> 4) & 0x0F0F0F0Fu) | spread4((qh >> (16 + 4 * i)) & 0xFu));
        }
    }
    __device__ __forceinline__ float dot(const act_unit & u) const {
        int sumi = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) sumi = __dp4a(v[i], u.a[i], sumi);
        return TYPE == B200_TYPE_Q5_0 ? d * (u.d * (float) sumi - 16.0f * u.s) : d * u.d * (float) sumi + m * u.s;
    }
};
template <> struct w_unit<B200_TYPE_Q5_0> : w_unit_q5<B200_TYPE_Q5_0> {};
template <> struct w_unit<B200_TYPE_Q5_1> : w_unit_q5<B200_TYPE_Q5_1> {};

template <> struct w_unit<B200_TYPE_Q5_K> {
    int v[8]; float dsc, dmn;
    __device__ __forceinline__ void load(const uint8_t * row, int64_t ib) {
        const int64_t sb = ib >> 3;
        const int     j  = (int) (ib & 7);
        const uint32_t * bp = reinterpret_cast<const uint32_t *>(row + sb * SZ_Q5_K);     // 176-byte blocks: 4-byte aligned
        const float2 dm = __half22float2(*reinterpret_cast<const half2 *>(bp));
        int sc, mn;
        q4k_scale_min(j, __ldg(bp + 1), __ldg(bp + 2), __ldg(bp + 3), sc, mn);
        dsc = dm.x * (float) sc;
        dmn = dm.y * (float) mn;
        const uint32_t * qh = bp + 4;                  // 32 bytes: bit j of byte l = fifth bit of element l of sub-block j
        const uint32_t * q  = bp + 12 + 8 * (j >> 1);  // 32 bytes: the nibbles of sub-blocks 2c (low) and 2c + 1 (high)
        const int sh = 4 * (j & 1);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = (int) (((__ldg(q + i) >> sh) & 0x0F0F0F0Fu) | (((__ldg(qh + i) >> j) & 0x01010101u) << 4));
    }
    __device__ __forceinline__ float dot(const act_unit & u) const {
        int sumi = 0, sum8 = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) { sumi = __dp4a(v[i], u.a[i], sumi); sum8 = __dp4a(0x01010101, u.a[i], sum8); }
        return u.d * (dsc * (float) sumi - dmn * (float) sum8);
    }
};

// N4 — Q2_K / Q3_K / IQ4_NL / IQ4_XS / MXFP4 (ggml-common.h:204-209, 288-311, 437-450; dequantize_row_* ggml-quants.c:507-525, 899-930, 1243-1290,
// 2649-2688).  Unit ib of a K-quant row = elements [32 j, 32 j + 32) of super-block ib / 8 with j = ib % 8: 128-half n = j / 4, two-bit field
// r = j % 4 of the 32 bytes qs[32 n ..], and two 16-element scales (2 j, 2 j + 1).
template <> struct w_unit<B200_TYPE_Q2_K> {
    int v[8]; float d0, d1, m0, m1;
    __device__ __forceinline__ void load(const uint8_t * row, int64_t ib) {
        const int64_t sb = ib >> 3;
        const int     j  = (int) (ib & 7), n = j >> 2, r = j & 3;
        const uint32_t * bp = reinterpret_cast<const uint32_t *>(row + sb * SZ_Q2_K);     // 84-byte blocks: 4-byte aligned
        const uint32_t sc = (__ldg(bp + 2 * n + (r >> 1)) >> (16 * (r & 1))) & 0xFFFFu;    // scales[8 n + 2 r], [.. + 1]: low nibble scale, high nibble minimum
        const float2 dm = __half22float2(*reinterpret_cast<const half2 *>(bp + 20));
        d0 = dm.x * (float) (sc & 0xF);        m0 = dm.y * (float) ((sc >> 4) & 0xF);
        d1 = dm.x * (float) ((sc >> 8) & 0xF); m1 = dm.y * (float) (sc >> 12);
        const uint32_t * q = bp + 4 + 8 * n;
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = (int) ((__ldg(q + i) >> (2 * r)) & 0x03030303u);
    }
    __device__ __forceinline__ float dot(const act_unit & u) const {
        int s0 = 0, s1 = 0, a0 = 0, a1 = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            s0 = __dp4a(v[i], u.a[i], s0);         a0 = __dp4a(0x01010101, u.a[i], a0);
            s1 = __dp4a(v[i + 4], u.a[i + 4], s1); a1 = __dp4a(0x01010101, u.a[i + 4], a1);
        }
        return u.d * (d0 * (float) s0 + d1 * (float) s1 - m0 * (float) a0 - m1 * (float) a1);
    }
};

template <> struct w_unit<B200_TYPE_Q3_K> {
    int v[8]; float d0, d1;
    __device__ __forceinline__ void load(const uint8_t * row, int64_t ib) {
        const int64_t sb = ib >> 3;
        const int     j  = (int) (ib & 7), n = j >> 2, r = j & 3;
        const uint8_t * bp = row + sb * SZ_Q3_K;                                          // 110-byte blocks: 2-byte aligned only
        uint32_t hw[8], lw[8];
        ldw_b2<8>(bp, hw);                       // hmask: bit j of byte l = third bit of element l of unit j (set: 0, clear: -4)
        ldw_b2<8>(bp + 32 + 32 * n, lw);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = (int) __vsubss4(((lw[i] >> (2 * r)) & 0x03030303u) | (((hw[i] >> j) & 0x01010101u) << 2), 0x04040404u);
        // sixteen 6-bit scales in 12 bytes: low nibbles of bytes 0..7 = scales 0..7, high nibbles = scales 8..15, bytes 8..11 hold the top two bits
        const uint8_t * raw = bp + 96;
        const int is = 2 * j, sh = 2 * (is >> 2);
        const int b0 = __ldg(raw + (is & 7)), b1 = __ldg(raw + (is & 7) + 1);
        const int l0 = j < 4 ? (b0 & 0xF) : (b0 >> 4), l1 = j < 4 ? (b1 & 0xF) : (b1 >> 4);
        const int h0 = (__ldg(raw + 8 + (is & 3)) >> sh) & 3, h1 = (__ldg(raw + 9 + (is & 3)) >> sh) & 3;
        const float d = ld_half(bp + 108);
        d0 = d * (float) ((l0 | (h0 << 4)) - 32);
        d1 = d * (float) ((l1 | (h1 << 4)) - 32);
    }
    __device__ __forceinline__ float dot(const act_unit & u) const {
        int s0 = 0, s1 = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) { s0 = __dp4a(v[i], u.a[i], s0); s1 = __dp4a(v[i + 4], u.a[i + 4], s1); }
        return u.d * (d0 * (float) s0 + d1 * (float) s1);
    }
};

// sixteen int8 code-book values (t0..t3 = entries 0-3, 4-7, 8-11, 12-15) looked up for the four 4-bit indices in the bytes of x
__host__ __device__ constexpr uint32_t pack_s8(int a, int b, int c, int d) {
    return (uint32_t) (a & 0xFF) | ((uint32_t) (b & 0xFF) << 8) | ((uint32_t) (c & 0xFF) << 16) | ((uint32_t) (d & 0xFF) << 24);
}
__device__ __forceinline__ int tab16(const uint32_t x, const uint32_t t0, const uint32_t t1, const uint32_t t2, const uint32_t t3) {
    const uint32_t y   = x & 0x07070707u;
    const uint32_t sel = __byte_perm(y | (y >> 4), 0u, 0x4420);            // the four 3-bit indices as PRMT selector nibbles
    const uint32_t lo  = __byte_perm(t0, t1, sel), hi = __byte_perm(t2, t3, sel);
    const uint32_t top = ((x >> 3) & 0x01010101u) * 0xFFu;                  // 0xFF where the index is 8..15
    return (int) ((hi & top) | (lo & ~top));
}
// kvalues_iq4nl / kvalues_mxfp4 (ggml-common.h:1110-1118)
#define IQ4NL_TAB pack_s8(-127, -104, -83, -65), pack_s8(-49, -35, -22, -10), pack_s8(1, 13, 25, 38), pack_s8(53, 69, 89, 113)
#define MXFP4_TAB pack_s8(0, 1, 2, 3), pack_s8(4, 6, 8, 12), pack_s8(0, -1, -2, -3), pack_s8(-4, -6, -8, -12)

template <> struct w_unit<B200_TYPE_IQ4_NL> {
    int lo[4], hi[4]; float d;
    __device__ __forceinline__ void load(const uint8_t * row, int64_t ib) {
        const uint8_t * bp = row + ib * SZ_IQ4_NL;
        d = ld_half(bp);
        uint32_t q[4];
        ldw_b2<4>(bp + 2, q);
#pragma unroll
        for (int i = 0; i < 4; ++i) { lo[i] = tab16(q[i] & 0x0F0F0F0Fu, IQ4NL_TAB); hi[i] = tab16((q[i] >> 4) & 0x0F0F0F0Fu, IQ4NL_TAB); }
    }
    __device__ __forceinline__ float dot(const act_unit & u) const {
        int sumi = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) { sumi = __dp4a(lo[i], u.a[i], sumi); sumi = __dp4a(hi[i], u.a[i + 4], sumi); }
        return d * u.d * (float) sumi;
    }
};

template <> struct w_unit<B200_TYPE_IQ4_XS> {
    int lo[4], hi[4]; float d;
    __device__ __forceinline__ void load(const uint8_t * row, int64_t ib) {
        const int64_t sb = ib >> 3;
        const int     j  = (int) (ib & 7);
        const uint32_t * bp = reinterpret_cast<const uint32_t *>(row + sb * SZ_IQ4_XS);   // 136-byte blocks: 8-byte aligned
        const uint32_t w0 = __ldg(bp), sl = __ldg(bp + 1);                                // d | scales_h << 16, scales_l[4]
        const int ls = (int) (((sl >> (4 * j)) & 0xFu) | ((((w0 >> 16) >> (2 * j)) & 3u) << 4));
        d = __half2float(__ushort_as_half((unsigned short) (w0 & 0xFFFFu))) * (float) (ls - 32);
        const uint32_t * q = bp + 2 + 4 * j;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t w = __ldg(q + i);
            lo[i] = tab16(w & 0x0F0F0F0Fu, IQ4NL_TAB); hi[i] = tab16((w >> 4) & 0x0F0F0F0Fu, IQ4NL_TAB);
        }
    }
    __device__ __forceinline__ float dot(const act_unit & u) const {
        int sumi = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) { sumi = __dp4a(lo[i], u.a[i], sumi); sumi = __dp4a(hi[i], u.a[i + 4], sumi); }
        return d * u.d * (float) sumi;
    }
};

// half of 2^(e - 127) (the code book holds the e2m1 values doubled; ggml-impl.h:477-495)
__device__ __forceinline__ float e8m0_half(const uint32_t e) { return __uint_as_float(e < 2 ? 0x00200000u << e : (e - 1) << 23); }

template <> struct w_unit<B200_TYPE_MXFP4> {
    int lo[4], hi[4]; float d;
    __device__ __forceinline__ void load(const uint8_t * row, int64_t ib) {
        const uint8_t * bp = row + ib * SZ_MXFP4;                                         // 17-byte blocks: no alignment at all
        const uint32_t * a = reinterpret_cast<const uint32_t *>(reinterpret_cast<uintptr_t>(bp) & ~(uintptr_t) 3);
        const uint32_t  sh = ((uint32_t) reinterpret_cast<uintptr_t>(bp) & 3u) * 8u;
        uint32_t w[6];
#pragma unroll
        for (int i = 0; i < 6; ++i) w[i] = __ldg(a + i);                                   // (up to 7 bytes past the block: rows are padded)
        d = e8m0_half(__funnelshift_r(w[0], w[1], sh) & 0xFFu);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t q = __funnelshift_r(__funnelshift_r(w[i], w[i + 1], sh), __funnelshift_r(w[i + 1], w[i + 2], sh), 8);
            lo[i] = tab16(q & 0x0F0F0F0Fu, MXFP4_TAB); hi[i] = tab16((q >> 4) & 0x0F0F0F0Fu, MXFP4_TAB);
        }
    }
    __device__ __forceinline__ float dot(const act_unit & u) const {
        int sumi = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) { sumi = __dp4a(lo[i], u.a[i], sumi); sumi = __dp4a(hi[i], u.a[i + 4], sumi); }
        return d * u.d * (float) sumi;
    }
};

// UE4M3 scale (unsigned, bias 7, 3 mantissa bits), halved like the code book is doubled (ggml-impl.h:500-516)
__device__ __forceinline__ float ue4m3_half(const uint32_t x) {
    if (x == 0 || x == 0x7F) return 0.0f;
    const uint32_t e = (x >> 3) & 0xFu, man = x & 7u;
    return e == 0 ? (float) man * 0.0009765625f /* 2^-10 */ : __uint_as_float(((e + 119u) << 23) | (man << 20));
}

// NVFP4 (ggml-common.h:211-217, dequantize_row_nvfp4 ggml-quants.c:527-550): unit ib = sub-blocks 2 h, 2 h + 1 of block ib / 2 (h = ib % 2); a
// sub-block's 8 bytes hold elements 0..7 in their low nibbles and 8..15 in their high nibbles
template <> struct w_unit<B200_TYPE_NVFP4> {
    int v[8]; float d0, d1;
    __device__ __forceinline__ void load(const uint8_t * row, int64_t ib) {
        const uint32_t * bp = reinterpret_cast<const uint32_t *>(row + (ib >> 1) * SZ_NVFP4);   // 36-byte blocks: 4-byte aligned
        const int h = (int) (ib & 1);
        const uint32_t sc = __ldg(bp) >> (16 * h);
        d0 = ue4m3_half(sc & 0xFFu); d1 = ue4m3_half((sc >> 8) & 0xFFu);
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            const uint32_t w0 = __ldg(bp + 1 + 4 * h + 2 * s), w1 = __ldg(bp + 2 + 4 * h + 2 * s);
            v[4 * s]     = tab16(w0 & 0x0F0F0F0Fu, MXFP4_TAB);        v[4 * s + 1] = tab16(w1 & 0x0F0F0F0Fu, MXFP4_TAB);
            v[4 * s + 2] = tab16((w0 >> 4) & 0x0F0F0F0Fu, MXFP4_TAB); v[4 * s + 3] = tab16((w1 >> 4) & 0x0F0F0F0Fu, MXFP4_TAB);
        }
    }
    __device__ __forceinline__ float dot(const act_unit & u) const {
        int s0 = 0, s1 = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) { s0 = __dp4a(v[i], u.a[i], s0); s1 = __dp4a(v[i + 4], u.a[i + 4], s1); }
        return u.d * (d0 * (float) s0 + d1 * (float) s1);
    }
};

constexpr int GENERIC_WARPS = 4;

template <int TYPE, int NC>
__global__ void __launch_bounds__(GENERIC_WARPS * 32)
gemv_generic(const mm_problem P) {
    mm_cols<NC> C;
    if (!mm_resolve<NC>(P, blockIdx.y, C)) return;
    const int     lane = threadIdx.x & 31;
    const int64_t row  = (int64_t) blockIdx.x * GENERIC_WARPS + (threadIdx.x >> 5);
    if (row >= P.m) return;                                   // whole warp leaves together
    const uint8_t * wrow = C.W + row * P.nb01;
    const int64_t   nu   = P.k >> 5;
    float acc[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) acc[c] = 0.0f;
    for (int64_t ib = lane; ib < nu; ib += 32) {
        w_unit<TYPE> w;
        w.load(wrow, ib);
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const act_unit u = load_act(P.act, C.plane_col[c], ib);
            acc[c] += w.dot(u);
        }
    }
#pragma unroll
    for (int c = 0; c < NC; ++c) {
        const float s = warp_sum(acc[c]);
        if (lane == 0 && c < C.nc) mm_store(P, C.y[c], row, s);
    }
}

// N2 — MUL_MAT (+ bias) (+ gate MUL_MAT (+ bias) + GLU) for one activation column per batch in one launch (the reference fuses the same
// pattern into its mul_mat_vec_q kernel: ggml-cuda/mmvq.cu:440-477, chosen by ggml_cuda_should_fuse_mul_mat_vec_q ggml-cuda.cu:2499-2535).
// One warp per output row walks the row of the up matrix and, if there is a gate, the same row of the gate matrix against ONE
// fetch of the quantized activations; the epilogue adds the biases and applies the activation, so neither product goes to memory.
struct mm_fuse {
    const uint8_t * Wg;                   // gate matrix (same type, shape and strides as P.W) or NULL
    const float * bias_up, * bias_gate;   // f32 [m, 1, ne12, ne13] or NULL
    int64_t bu_nb2, bu_nb3, bg_nb2, bg_nb3;
    int glu_op;
};

template <int TYPE>
__global__ void __launch_bounds__(GENERIC_WARPS * 32)
gemv_fused(const mm_problem P, const mm_fuse F) {
    mm_cols<1> C;
    if (!mm_resolve<1>(P, blockIdx.y, C)) return;
    const int     lane = threadIdx.x & 31;
    const int64_t row  = (int64_t) blockIdx.x * GENERIC_WARPS + (threadIdx.x >> 5);
    if (row >= P.m) return;                                   // whole warp leaves together
    const uint8_t * urow = C.W + row * P.nb01;
    const uint8_t * grow = F.Wg ? F.Wg + (C.W - P.W) + row * P.nb01 : nullptr;
    const int64_t   nu   = P.k >> 5;
    float au = 0.0f, ag = 0.0f;
    for (int64_t ib = lane; ib < nu; ib += 32) {
        const act_unit u = load_act(P.act, C.plane_col[0], ib);
        w_unit<TYPE> w;
        w.load(urow, ib);
        au += w.dot(u);
        if (grow) {
            w_unit<TYPE> g;
            g.load(grow, ib);
            ag += g.dot(u);
        }
    }
    au = warp_sum(au);
    if (grow) ag = warp_sum(ag);
    if (lane == 0) {
        const int64_t batch = blockIdx.y, i12 = batch % P.ne12, i13 = batch / P.ne12;
        if (F.bias_up) au += *reinterpret_cast<const float *>(reinterpret_cast<const char *>(F.bias_up) + row * 4 + i12 * F.bu_nb2 + i13 * F.bu_nb3);
        if (grow) {
            if (F.bias_gate) ag += *reinterpret_cast<const float *>(reinterpret_cast<const char *>(F.bias_gate) + row * 4 + i12 * F.bg_nb2 + i13 * F.bg_nb3);
            au = glu_value(F.glu_op, ag, au, 0.0f, 0.0f);
        }
        C.y[0][row] = au;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Q4_K fast path
// ------------------------------------------------------------------------------------------------------------------
// Shared memory per CTA: NC columns of { int8 qs[k] in swizzled 16-byte chunks ; float2 (d, d*sum) [k/32] }.
//
// A lane's task = (superblock sb, 64-group g in 0..3): the 32 qs bytes of group g hold sub-block 2g in their low nibbles
// and sub-block 2g+1 in their high nibbles (dequantize_row_q4_K, ggml/src/ggml-quants.c:1467-1489).  4 consecutive lanes
// cover one 144-byte superblock (header re-read by the 4 lanes: one sector), a warp covers 8 superblocks = 1152
// contiguous bytes per step with three LDG.128 per lane.  High nibbles are NOT shifted down: (q & 0xF0F0F0F0) is fed to
// an unsigned x signed dp4a and the factor 1/16 is folded into the float scale (exact).
//
// Activation chunk t (16 int8) of a column lives at 16-byte slot  t ^ (((t>>3)&1)<<1) ^ ((t>>4)&1): for a lane the four
// chunks of its 64-byte activation group are permuted by a lane-constant x in 0..3, and the 8 lanes of every LDS.128
// phase hit 8 distinct 16-byte bank groups.
__device__ __forceinline__ int act_slot(int t) { return t ^ (((t >> 3) & 1) << 1) ^ ((t >> 4) & 1); }

__device__ __forceinline__ int dp4a_us(uint32_t a_unsigned, int b_signed, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a_unsigned), "r"(b_signed), "r"(c));
    return d;
}

struct q4k_frag { uint4 hdr, q0, q1; };

// whole-row prefetch into L2 (address and size are multiples of 16: a Q4_K row is 144-byte superblocks)
__device__ __forceinline__ void l2_prefetch(const void * p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(p), "r"(bytes) : "memory");
}

__device__ __forceinline__ void q4k_load(q4k_frag & f, const uint8_t * wrow, const int sb, const int g) {
    const uint8_t * bp = wrow + (size_t) sb * SZ_Q4_K;
    f.hdr = ld_stream_u4(bp);
    f.q0  = ld_stream_u4(bp + 16 + 32 * g);
    f.q1  = ld_stream_u4(bp + 32 + 32 * g);
}

template <int NC>
__device__ __forceinline__ void q4k_task(const q4k_frag & f, const int sb, const int g, const int x, const uint8_t * s_qs,
                                         const float2 * s_ds, const int k, float (&acc)[NC]) {
    // (sc, m) of sub-blocks 2g and 2g+1 as byte pairs — get_scale_min_k4, ggml/src/ggml-quants.c:818-825
    const int sh = 16 * (g & 1);
    const uint32_t a = f.hdr.y >> sh, b = f.hdr.z >> sh, c = f.hdr.w >> sh;
    uint32_t scp, mnp;
    if (g < 2) {
        scp = a & 0x3f3f;
        mnp = b & 0x3f3f;
    } else {
        scp = (c & 0x0f0f) | ((a & 0xc0c0) >> 2);
        mnp = ((c >> 4) & 0x0f0f) | ((b & 0xc0c0) >> 2);
    }
    const float2 dm  = __half22float2(*reinterpret_cast<const half2 *>(&f.hdr.x));
    const float dsc0 = dm.x * (float) (scp & 0xff);
    const float dsc1 = dm.x * (float) (scp >> 8) * 0.0625f;          // high nibbles enter the dot product as 16*q
    const float dmn0 = dm.y * (float) (mnp & 0xff);
    const float dmn1 = dm.y * (float) (mnp >> 8);
    const uint32_t m_lo = 0x0F0F0F0Fu, m_hi = 0xF0F0F0F0u;
    const int ib0 = sb * 8 + 2 * g;                                   // activation block of sub-block 2g
    const int off = (sb * 16 + 4 * g) * 16;                           // byte offset of the lane's 64-byte activation group
#pragma unroll
    for (int col = 0; col < NC; ++col) {
        const uint8_t * cq = s_qs + (size_t) col * k + off;
        const uint4 a0 = *reinterpret_cast<const uint4 *>(cq + 16 * (0 ^ x));
        const uint4 a1 = *reinterpret_cast<const uint4 *>(cq + 16 * (1 ^ x));
        const uint4 a2 = *reinterpret_cast<const uint4 *>(cq + 16 * (2 ^ x));
        const uint4 a3 = *reinterpret_cast<const uint4 *>(cq + 16 * (3 ^ x));
        const float4 e = *reinterpret_cast<const float4 *>(s_ds + col * (k >> 5) + ib0);   // (d, d*sum) of blocks ib0, ib0+1
        int s0 = __dp4a((int) (f.q0.x & m_lo), (int) a0.x, 0);
        s0 = __dp4a((int) (f.q0.y & m_lo), (int) a0.y, s0); s0 = __dp4a((int) (f.q0.z & m_lo), (int) a0.z, s0); s0 = __dp4a((int) (f.q0.w & m_lo), (int) a0.w, s0);
        s0 = __dp4a((int) (f.q1.x & m_lo), (int) a1.x, s0); s0 = __dp4a((int) (f.q1.y & m_lo), (int) a1.y, s0);
        s0 = __dp4a((int) (f.q1.z & m_lo), (int) a1.z, s0); s0 = __dp4a((int) (f.q1.w & m_lo), (int) a1.w, s0);
        int s1 = dp4a_us(f.q0.x & m_hi, (int) a2.x, 0);
        s1 = dp4a_us(f.q0.y & m_hi, (int) a2.y, s1); s1 = dp4a_us(f.q0.z & m_hi, (int) a2.z, s1); s1 = dp4a_us(f.q0.w & m_hi, (int) a2.w, s1);
        s1 = dp4a_us(f.q1.x & m_hi, (int) a3.x, s1); s1 = dp4a_us(f.q1.y & m_hi, (int) a3.y, s1);
        s1 = dp4a_us(f.q1.z & m_hi, (int) a3.z, s1); s1 = dp4a_us(f.q1.w & m_hi, (int) a3.w, s1);
        float r = acc[col];
        r = fmaf(dsc0 * e.x, (float) s0, r);
        r = fmaf(dsc1 * e.z, (float) s1, r);
        r = fmaf(-dmn0, e.y, r);
        r = fmaf(-dmn1, e.w, r);
        acc[col] = r;
    }
}

// VARIANT: 0 = plain, 1 = cross-rank completion protocol (row split, one process per GPU), 2 = several matrices per launch,
//          3 = several matrices per launch + LL exchange of the results with all ranks (row split, one process per GPU)
// mm_resolve for the common case "one matrix, one column tile" (every decode matmul of a model): nothing to divide.  The 64-bit
// divisions of the general resolution cost about a microsecond on the way to the first weight request of a kernel that
// streams its matrix in two (measured: Llama-3-8B decode step 1.20 -> 1.11 ms).
// Why this is a separate function and not a branch inside mm_resolve (round-2 diagnosis of the round-1 conformance failure):
// with the shortcut INSIDE mm_resolve — one call site or two, Wm const or not — the nvcc 12.9 / ptxas build of gemv_q4k<1,32>
// computes the FIRST weight fragment of the early (pre-griddepcontrol) path from P.W read straight out of the constant bank
// (SASS at the join of the two resolve branches: LDCU.128 c[0x0][0x380] -> address = P.W + row * nb01 + sb * 144), i.e. it drops
// the batch offset (i12 / r2) * nb02 + (i13 / r3) * nb03 that the general branch computed; x and y pointers are right.  Every
// batched test-backend-ops case (bs != [1,1]) then multiplies batch 0's weights; with pdl = 0 (the late path) the same binary is
// correct, and an in-kernel printf of C.W shows the resolved pointer correct.  Kept apart, the compiler has nothing to fold.
template <int NC>
__device__ __forceinline__ bool mm_resolve_fast(const mm_problem & P, mm_cols<NC> & C) {
    if (!(P.mode == MM_PLAIN && gridDim.y == 1)) return mm_resolve<NC>(P, blockIdx.y, C);
    C.W  = P.W;
    C.nc = (int) (P.ne11 < NC ? P.ne11 : NC);
#pragma unroll
    for (int c = 0; c < NC; ++c) {
        const int64_t i11 = c < C.nc ? c : 0;
        C.plane_col[c] = i11;
        C.x[c] = reinterpret_cast<const float *>(reinterpret_cast<const char *>(P.X) + i11 * P.nb11);
        C.y[c] = P.Y + i11 * P.m;
    }
    return true;
}

template <int NC, int NWARPS, int MINB, int NBUF, int VARIANT>
__global__ void __launch_bounds__(NWARPS * 32, MINB)
gemv_q4k(const mm_problem P, const int flags) {
    constexpr bool SYNC = VARIANT == 1, LL = VARIANT == 3, MULTI = VARIANT == 2 || LL;
    // flags: < 0 launch-floor probe (ctx option "noop"); bit 0 = PDL launch (prefetch weights before the dependency wait);
    //        bit 1 = X / ids are NOT written by any kernel still in flight (abi.cu tracks that): the wait moves from
    //        "before reading X" to "before the first store", so this kernel's prologue and streaming overlap the previous
    //        kernel's tail, while completion stays transitive (nobody finishes before its predecessor).
    if (flags < 0) return;
    const int  prefetch_before_wait = flags & 1;
    const bool late_wait            = (flags & 2) != 0;
    extern __shared__ __align__(16) uint8_t smem[];
    const int k   = (int) P.k;
    const int nbk = k >> 5;
    const int nsb = k >> 8;
    uint8_t * s_qs = smem;
    float2 *  s_ds = reinterpret_cast<float2 *>(smem + (size_t) NC * k);

    const int     lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int     sbl  = lane >> 2, g = lane & 3;
    const int     x    = ((g >> 1) << 1) ^ (sbl & 1);
    // this CTA's contiguous share of the rows (32-bit arithmetic whenever it cannot overflow: no 64-bit divisions on the way
    // to the first weight request)
    const int64_t rlo  = P.m < (1 << 22) ? (int64_t) ((uint32_t) P.m * blockIdx.x / gridDim.x) : (P.m * (int64_t) blockIdx.x) / gridDim.x;
    const int64_t rhi  = P.m < (1 << 22) ? (int64_t) ((uint32_t) P.m * (blockIdx.x + 1) / gridDim.x) : (P.m * (int64_t) (blockIdx.x + 1)) / gridDim.x;

    // Weights are never produced by a kernel of this path, so this warp's first rows are requested before waiting on the
    // previous kernel (PDL) and before the activation prologue (first fragment into registers).  The main loop keeps the
    // register pipeline one fragment ahead (load next, then consume current).  NBUF=1 adds a bulk L2 prefetch two rows
    // ahead; measured slower on B200 (output head 65 us vs 54 us), kept only as a tuning switch.
    const int nit = (nsb + 7) >> 3;                       // steps per row; (row, it) are warp-uniform, sb differs per lane
    const uint32_t row_bytes = (uint32_t) nsb * SZ_Q4_K;
    int64_t  row = rlo + warp;
    int      it  = 0, sb = sbl;
    q4k_frag cur;
    const uint8_t * Wm = nullptr;
    auto wrow = [&](const int64_t r) -> const uint8_t * {
        const uint8_t * base = Wm;
        if (MULTI) base = r < P.m_end[0] ? Wm : (r < P.m_end[1] ? P.Wv[0] : P.Wv[1]);
        return base + r * P.nb01;
    };
    auto first_loads = [&]() {
        if (NBUF > 0 && lane == 0) {
            if (row < rhi)          l2_prefetch(wrow(row), row_bytes);
            if (row + NWARPS < rhi) l2_prefetch(wrow(row + NWARPS), row_bytes);
        }
        if (row < rhi && sb < nsb) q4k_load(cur, wrow(row), sb, g);
    };

    const bool early = prefetch_before_wait && (P.mode == MM_PLAIN || late_wait);
    mm_cols<NC> C;
    bool live = true;
    if (early) {
        live = mm_resolve_fast<NC>(P, C);
        Wm = C.W;
        if (live) first_loads();
    }

    // the previous kernel in the stream may have produced X or the ids (no-op unless launched with the PDL attribute)
    if (late_wait) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    else           asm volatile("griddepcontrol.wait;" ::: "memory");
    if (!early) {
        live = mm_resolve<NC>(P, blockIdx.y, C);
        Wm = C.W;
        if (live) first_loads();
    }
    if (!live) return;                                          // block-uniform
    // LL exchange: this launch's epoch = the last completed one (ll_state[0], advanced by the receiver kernel) + 1.  CTA 0
    // hands it to this launch's receiver through ll_state[1] — performed (fence) before this CTA's launch_dependents below,
    // and the receiver cannot become resident before every CTA of this grid has passed that point.
    uint32_t ll_e = 0;
    if (LL) {
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(ll_e) : "l"(P.ll_state) : "memory");
        asm volatile("add.u32 %0, %0, 1;" : "+r"(ll_e));       // consumes the load here: it has been performed before the trigger
        if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) {
            asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" :: "l"(P.ll_state + 1), "r"(ll_e) : "memory");
            __threadfence();
        }
    }
    if (SYNC && P.sync_flags) {
        // fused row-split all-gather: the activations may derive from the previous matmul's gathered output, which is complete
        // once every rank has published that launch's epoch (our own previous launch is complete: griddepcontrol.wait)
        if (threadIdx.x == 0) {
            const uint32_t e = *reinterpret_cast<volatile uint32_t *>(P.sync_state);
            const long long t0 = clock64();
            for (int src = 0; src < P.sync_world; ++src) {
                uint32_t seen;
                do {
                    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(P.sync_flags + src) : "memory");
                    if (clock64() - t0 > 6000000000LL) __trap();        // a lost rank must not hang the others
                } while ((int32_t) (seen - e) < 0);
            }
        }
        __syncthreads();
    }

    // ---- prologue: quantize the NC activation columns into shared memory (bit-exact Q8_1 values) ----
    {
        const int sub    = threadIdx.x & 7;
        const int total  = NC * nbk;
        const int groups = NWARPS * 4;
        const int rounds = (total + groups - 1) / groups;
        int blk = threadIdx.x >> 3;
        for (int r = 0; r < rounds; ++r, blk += groups) {
            const bool ok  = blk < total;
            const int  b   = ok ? blk : total - 1;
            const int  col = NC == 1 ? 0 : b / nbk;
            const int  ib  = b - col * nbk;
            const float * xc = C.x[0];
#pragma unroll
            for (int c = 1; c < NC; ++c) if (col == c) xc = C.x[c];
            const float4 v = __ldg(reinterpret_cast<const float4 *>(xc + ib * 32 + sub * 4));
            const q8_part qp = quantize_q8_1_part(v);
            if (ok) {
                const int t = ib * 2 + (sub >> 2);
                *reinterpret_cast<uint32_t *>(s_qs + (size_t) col * k + 16 * act_slot(t) + 4 * (sub & 3)) = qp.packed;
                if (sub == 0) s_ds[col * nbk + ib] = make_float2(qp.d, __fmul_rn((float) qp.sum, qp.d));
            }
        }
    }
    __syncthreads();
    if (!late_wait) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    bool waited = !late_wait;

    // ---- main loop: this CTA's contiguous share of the rows, one warp per row ----
    float acc[NC];
#pragma unroll
    for (int col = 0; col < NC; ++col) acc[col] = 0.0f;
    while (row < rhi) {
        int64_t nrow  = row;
        int     nit_i = it + 1;
        if (nit_i >= nit) { nrow = row + NWARPS; nit_i = 0; }
        const int nsbi = nit_i * 8 + sbl;
        q4k_frag nxt;
        if (nrow < rhi && nsbi < nsb) q4k_load(nxt, wrow(nrow), nsbi, g);
        if (NBUF > 0 && it == 0 && lane == 0 && row + 2 * NWARPS < rhi) l2_prefetch(wrow(row + 2 * NWARPS), row_bytes);
        if (sb < nsb) q4k_task<NC>(cur, sb, g, x, s_qs, s_ds, k, acc);
        if (nrow != row) {
            if (!waited) { asm volatile("griddepcontrol.wait;" ::: "memory"); waited = true; }
#pragma unroll
            for (int col = 0; col < NC; ++col) {
                const float s = warp_sum(acc[col]);
                if (LL) {
                    if (lane == 0) {
                        const int64_t slot = row + (row < P.m_end[0] ? P.ll_off[0] : (row < P.m_end[1] ? P.ll_off[1] : P.ll_off[2])) +
                                             (int64_t) (ll_e & 1) * P.ll_cap;
#pragma unroll 1
                        for (int p = 0; p < P.ll_world; ++p) {
                            asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" :: "l"(P.ll_dst[p] + slot), "r"(__float_as_uint(s)), "r"(ll_e) : "memory");
                        }
                    }
                } else if (MULTI) {
                    if (lane == 0) (row < P.m_end[0] ? C.y[0] : (row < P.m_end[1] ? P.Yv[0] : P.Yv[1]))[row] = s;
                } else if (lane == 0 && col < C.nc) {
                    mm_store(P, C.y[col], row, s);
                }
                acc[col] = 0.0f;
            }
        }
        cur = nxt;
        row = nrow;
        it  = nit_i;
        sb  = nsbi;
    }
    if (SYNC && P.sync_flags) {
        // publish: every CTA makes its peer stores visible system-wide, the last one to finish bumps the epoch and writes it
        // into flag[my rank] of every rank
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence_system();                            // cumulative: covers the stores of the whole CTA (bar.sync above)
            const unsigned total = gridDim.x * gridDim.y;
            if (atomicAdd(P.sync_state + 1, 1u) == total - 1) {
                P.sync_state[1] = 0;
                const uint32_t e = P.sync_state[0] + 1;
                P.sync_state[0] = e;
                __threadfence_system();
                for (int p = 0; p < P.sync_world; ++p) {
                    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(P.peer_flags[p] + P.sync_rank), "r"(e) : "memory");
                }
            }
        }
    }
}

template <int TYPE>
int launch_generic_t(b200_ctx * ctx, const mm_problem & P, int nc, int64_t tiles) {
    const dim3 grid((unsigned) ((P.m + GENERIC_WARPS - 1) / GENERIC_WARPS), (unsigned) tiles);
    const dim3 block(GENERIC_WARPS * 32);
    switch (nc) {
        case 1: gemv_generic<TYPE, 1><<<grid, block, 0, ctx->stream>>>(P); break;
        case 2: gemv_generic<TYPE, 2><<<grid, block, 0, ctx->stream>>>(P); break;
        case 3: gemv_generic<TYPE, 3><<<grid, block, 0, ctx->stream>>>(P); break;
        case 4: gemv_generic<TYPE, 4><<<grid, block, 0, ctx->stream>>>(P); break;
        case 5: gemv_generic<TYPE, 5><<<grid, block, 0, ctx->stream>>>(P); break;
        case 6: gemv_generic<TYPE, 6><<<grid, block, 0, ctx->stream>>>(P); break;
        case 7: gemv_generic<TYPE, 7><<<grid, block, 0, ctx->stream>>>(P); break;
        case 8: gemv_generic<TYPE, 8><<<grid, block, 0, ctx->stream>>>(P); break;
        default: b200_set_error("gemv_generic: nc=%d out of range", nc); return B200_E_ARG;
    }
    ctx->launches++;
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
}

// receiver side of the LL exchange: wait until every entry of the slot carries `epoch`, then store the values as plain f32
// vectors (one per matrix of the fused launch).  A rank that never delivers must not hang the others: trap after ~3 s.
struct ll_unpack_args {
    const uint2 * buf;                // both slots of this rank's exchange buffer
    int64_t       cap;
    uint32_t *    state;              // [0] epoch of the last completed exchange (written here), [1] epoch of the launch in flight (by its GEMV)
    int64_t       total, end[2];      // entries; matrix boundaries
    float *       Yv[3];              // virtual bases: entry i of matrix j is stored to Yv[j][i]
};

__global__ void __launch_bounds__(256)
ll_unpack_kernel(const ll_unpack_args a) {
    // let the next kernel of the stream (the next GEMV) become resident and prefetch its weights while this one polls; its
    // griddepcontrol.wait still returns only when this grid has completed
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int64_t i  = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t ii = i < a.total ? i : 0;
    // epoch and both candidate slots are requested together (one memory round trip when the values are already there)
    uint32_t epoch, v0, e0, v1, e1;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(epoch) : "l"(a.state + 1) : "memory");
    asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v0), "=r"(e0) : "l"(a.buf + ii) : "memory");
    asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v1), "=r"(e1) : "l"(a.buf + a.cap + ii) : "memory");
    // the next GEMV reads state[0] only after this whole grid has completed, and this launch's GEMV has read it long ago
    if (i == 0) asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" :: "l"(a.state), "r"(epoch) : "memory");
    if (i < a.total) {
        const uint2 * src = a.buf + (int64_t) (epoch & 1) * a.cap + i;
        uint32_t v = (epoch & 1) ? v1 : v0, e = (epoch & 1) ? e1 : e0;
        if (e != epoch) {
            const long long t0 = clock64();
            do {
                __nanosleep(32);
                asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v), "=r"(e) : "l"(src) : "memory");
                if (clock64() - t0 > 6000000000LL) __trap();
            } while (e != epoch);
        }
        (i < a.end[0] ? a.Yv[0] : (i < a.end[1] ? a.Yv[1] : a.Yv[2]))[i] = __uint_as_float(v);
    }
}

template <int NC, int NWARPS, int MINB, int NBUF, int VARIANT = 0>
int launch_q4k_t(b200_ctx * ctx, const mm_problem & P, int64_t tiles, size_t smem, bool late_wait) {
    if (VARIANT == 0 && (P.sync_flags || P.n_mats > 1 || P.ll_world > 0)) {
        // the cross-rank completion protocol and the several-matrices form live in their own instantiations so that the plain
        // kernel pays nothing for them
        if (NC == 1 && NWARPS == 32 && MINB == 1 && NBUF == 0 && P.ll_world > 0 && !P.sync_flags) {
            return launch_q4k_t<1, 32, 1, 0, 3>(ctx, P, tiles, smem, late_wait);
        }
        if (NC == 1 && NWARPS == 32 && MINB == 1 && NBUF == 0 && !(P.sync_flags && P.n_mats > 1)) {
            return P.sync_flags ? launch_q4k_t<1, 32, 1, 0, 1>(ctx, P, tiles, smem, late_wait) : launch_q4k_t<1, 32, 1, 0, 2>(ctx, P, tiles, smem, late_wait);
        }
        b200_set_error("gemv_q4k: this variant is built for n = 1 with the default CTA shape");
        return B200_E_UNSUPPORTED;
    }
    auto kern = gemv_q4k<NC, NWARPS, MINB, NBUF, VARIANT>;
    static std::atomic<bool> attr_done[16];   // per device (several backends of one device may launch from different threads)
    if (!attr_done[ctx->device & 15]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) ctx->smem_optin));
        attr_done[ctx->device & 15] = true;
    }
    // persistent: the SMs are shared evenly by the tiles (batches / (token, slot) pairs)
    int64_t ctas = (int64_t) ctx->sm_count * MINB / tiles;
    // Small matrices deliberately do NOT spread over every SM: one row per warp is the useful maximum, and the SMs left free
    // let the next GEMV of the stream start (PDL + deferred wait) while this one is still running — measured 2.4 us per
    // 1024x4096 matmul with 32 CTAs against 4.1 us with 148.
    // rows per warp that count as "useful" (tuning switch gemv_rpw, default 1; 0 = two rows per warp for 2048 <= m <= gemv_rpw_max_m).
    // Two rows per warp look better shape by shape (same shape back to back in a graph: 4096x4096 3.92 -> 3.31 us, 4096x14336
    // 8.91 -> 8.25 us) but make the real step SLOWER (1.110 -> 1.200 ms): half the CTAs live twice as long and delay the large
    // kernel that follows.  Per-shape timings are not additive; the step is what counts.
    const int64_t rpw = ctx->gemv_rpw >= 1 ? ctx->gemv_rpw : (P.m >= 2048 && P.m <= ctx->gemv_rpw_max_m ? 2 : 1);
    const int64_t max_useful = (P.m + NWARPS * rpw - 1) / (NWARPS * rpw);
    if (ctas > max_useful) ctas = max_useful;
    if (ctas < 1) ctas = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim          = dim3((unsigned) ctas, (unsigned) tiles);
    cfg.blockDim         = dim3(NWARPS * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream           = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs    = attr;
    cfg.numAttrs = ctx->pdl ? 1 : 0;
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, P, ctx->noop ? -1 : (ctx->pdl ? 1 | (late_wait ? 2 : 0) : 0)));
    ctx->launches++;
    return B200_OK;
}

}  // namespace

// ---- host entry points used by mul_mat.cu -------------------------------------------------------------------------

// pre-quantized activations in P.act; nc = columns per tile (1..8); tiles = gridDim.y
int b200_launch_gemv_fused(b200_ctx * ctx, int type, const mm_problem & P, const uint8_t * Wg, const float * bias_up, const int64_t * bu_nb,
                           const float * bias_gate, const int64_t * bg_nb, int glu_op, int64_t batches) {
    mm_fuse F = {};
    F.Wg = Wg; F.bias_up = bias_up; F.bias_gate = bias_gate; F.glu_op = glu_op;
    if (bias_up)   { F.bu_nb2 = bu_nb[0]; F.bu_nb3 = bu_nb[1]; }
    if (bias_gate) { F.bg_nb2 = bg_nb[0]; F.bg_nb3 = bg_nb[1]; }
    const dim3 grid((unsigned) ((P.m + GENERIC_WARPS - 1) / GENERIC_WARPS), (unsigned) batches);
    switch (type) {
        case B200_TYPE_Q4_0: gemv_fused<B200_TYPE_Q4_0><<<grid, GENERIC_WARPS * 32, 0, ctx->stream>>>(P, F); break;
        case B200_TYPE_Q8_0: gemv_fused<B200_TYPE_Q8_0><<<grid, GENERIC_WARPS * 32, 0, ctx->stream>>>(P, F); break;
        case B200_TYPE_Q4_K: gemv_fused<B200_TYPE_Q4_K><<<grid, GENERIC_WARPS * 32, 0, ctx->stream>>>(P, F); break;
        case B200_TYPE_Q6_K: gemv_fused<B200_TYPE_Q6_K><<<grid, GENERIC_WARPS * 32, 0, ctx->stream>>>(P, F); break;
        case B200_TYPE_Q4_1: gemv_fused<B200_TYPE_Q4_1><<<grid, GENERIC_WARPS * 32, 0, ctx->stream>>>(P, F); break;
        case B200_TYPE_Q5_0: gemv_fused<B200_TYPE_Q5_0><<<grid, GENERIC_WARPS * 32, 0, ctx->stream>>>(P, F); break;
        case B200_TYPE_Q5_1: gemv_fused<B200_TYPE_Q5_1><<<grid, GENERIC_WARPS * 32, 0, ctx->stream>>>(P, F); break;
        case B200_TYPE_Q5_K: gemv_fused<B200_TYPE_Q5_K><<<grid, GENERIC_WARPS * 32, 0, ctx->stream>>>(P, F); break;
        case B200_TYPE_Q2_K: gemv_fused<B200_TYPE_Q2_K><<<grid, GENERIC_WARPS * 32, 0, ctx->stream>>>(P, F); break;
        case B200_TYPE_Q3_K: gemv_fused<B200_TYPE_Q3_K><<<grid, GENERIC_WARPS * 32, 0, ctx->stream>>>(P, F); break;
        case B200_TYPE_IQ4_NL: gemv_fused<B200_TYPE_IQ4_NL><<<grid, GENERIC_WARPS * 32, 0, ctx->stream>>>(P, F); break;
        case B200_TYPE_IQ4_XS: gemv_fused<B200_TYPE_IQ4_XS><<<grid, GENERIC_WARPS * 32, 0, ctx->stream>>>(P, F); break;
        case B200_TYPE_MXFP4: gemv_fused<B200_TYPE_MXFP4><<<grid, GENERIC_WARPS * 32, 0, ctx->stream>>>(P, F); break;
        case B200_TYPE_NVFP4: gemv_fused<B200_TYPE_NVFP4><<<grid, GENERIC_WARPS * 32, 0, ctx->stream>>>(P, F); break;
        default: b200_set_error("fused GEMV: unsupported weight type %d", type); return B200_E_UNSUPPORTED;
    }
    B200_CUDA_TRY(cudaGetLastError());
    ctx->launches++;
    return B200_OK;
}

int b200_launch_gemv_generic(b200_ctx * ctx, int type, const mm_problem & P, int nc, int64_t tiles) {
    if (tiles <= 0 || P.m <= 0) return B200_OK;
    if (tiles > 65535) { b200_set_error("gemv: %lld column tiles exceed gridDim.y", (long long) tiles); return B200_E_ARG; }
    switch (type) {
        case B200_TYPE_Q4_0: return launch_generic_t<B200_TYPE_Q4_0>(ctx, P, nc, tiles);
        case B200_TYPE_Q8_0: return launch_generic_t<B200_TYPE_Q8_0>(ctx, P, nc, tiles);
        case B200_TYPE_Q4_K: return launch_generic_t<B200_TYPE_Q4_K>(ctx, P, nc, tiles);
        case B200_TYPE_Q6_K: return launch_generic_t<B200_TYPE_Q6_K>(ctx, P, nc, tiles);
        case B200_TYPE_Q4_1: return launch_generic_t<B200_TYPE_Q4_1>(ctx, P, nc, tiles);
        case B200_TYPE_Q5_0: return launch_generic_t<B200_TYPE_Q5_0>(ctx, P, nc, tiles);
        case B200_TYPE_Q5_1: return launch_generic_t<B200_TYPE_Q5_1>(ctx, P, nc, tiles);
        case B200_TYPE_Q5_K: return launch_generic_t<B200_TYPE_Q5_K>(ctx, P, nc, tiles);
        case B200_TYPE_Q2_K: return launch_generic_t<B200_TYPE_Q2_K>(ctx, P, nc, tiles);
        case B200_TYPE_Q3_K: return launch_generic_t<B200_TYPE_Q3_K>(ctx, P, nc, tiles);
        case B200_TYPE_IQ4_NL: return launch_generic_t<B200_TYPE_IQ4_NL>(ctx, P, nc, tiles);
        case B200_TYPE_IQ4_XS: return launch_generic_t<B200_TYPE_IQ4_XS>(ctx, P, nc, tiles);
        case B200_TYPE_MXFP4: return launch_generic_t<B200_TYPE_MXFP4>(ctx, P, nc, tiles);
        case B200_TYPE_NVFP4: return launch_generic_t<B200_TYPE_NVFP4>(ctx, P, nc, tiles);
    }
    b200_set_error("gemv: unsupported weight type %d", type);
    return B200_E_UNSUPPORTED;
}

// LL exchange, receiver side: entries [0, total) of `slot` -> n_mats plain vectors of m_full[j] floats
int b200_launch_ll_unpack(b200_ctx * ctx, const uint2 * buf, int64_t cap, uint32_t * state, int n_mats, const int64_t * m_full, float * const * Y) {
    ll_unpack_args a = {};
    a.buf = buf; a.cap = cap; a.state = state;
    int64_t start = 0;
    for (int j = 0; j < 3; ++j) {
        const int jj = j < n_mats ? j : n_mats - 1;
        if (j < n_mats) { a.Yv[j] = Y[j] - start; start += m_full[j]; } else a.Yv[j] = a.Yv[jj];
        if (j < 2) a.end[j] = j < n_mats ? start : INT64_MAX;
    }
    a.total = start;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim  = dim3((unsigned) ((a.total + 255) / 256));
    cfg.blockDim = dim3(256);
    cfg.stream   = ctx->stream;
    // programmatic launch: the receiver may become resident while the GEMV that feeds it is still running (it only polls)
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs    = attr;
    cfg.numAttrs = ctx->pdl ? 1 : 0;
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, ll_unpack_kernel, a));
    ctx->launches++;
    return B200_OK;
}

// shared memory the fused Q4_K kernel needs for nc columns of k values
size_t b200_gemv_q4k_smem(int nc, int64_t k) { return (size_t) nc * (size_t) (k + k / 4); }

// f32 activations in P.X (quantized in the kernel prologue); nc = columns of every tile (exact, 1..8)
int b200_launch_gemv_q4k(b200_ctx * ctx, const mm_problem & P, int nc, int64_t tiles, bool late_wait) {
    if (tiles <= 0 || (P.m <= 0 && P.ll_world == 0)) return B200_OK;    // (LL exchange: a rank without rows still announces the epoch)
    if (tiles > 65535) { b200_set_error("gemv_q4k: %lld tiles exceed gridDim.y", (long long) tiles); return B200_E_ARG; }
    const size_t smem = b200_gemv_q4k_smem(nc, P.k);
    if (smem > ctx->smem_optin) { b200_set_error("gemv_q4k: %zu B of shared memory needed", smem); return B200_E_ARG; }
    if (nc == 1) {
        switch (ctx->gemv_cfg) {
            case 1:  return launch_q4k_t<1, 16, 2, 0>(ctx, P, tiles, smem, late_wait);
            case 2:  return launch_q4k_t<1, 32, 1, 1>(ctx, P, tiles, smem, late_wait);      // + bulk L2 row prefetch (measured: slower)
            default: return launch_q4k_t<1, 32, 1, 0>(ctx, P, tiles, smem, late_wait);
        }
    }
    switch (nc) {
        case 2: return launch_q4k_t<2, 16, 1, 0>(ctx, P, tiles, smem, late_wait);
        case 3: return launch_q4k_t<3, 16, 1, 0>(ctx, P, tiles, smem, late_wait);
        case 4: return launch_q4k_t<4, 16, 1, 0>(ctx, P, tiles, smem, late_wait);
        case 5: return launch_q4k_t<5, 16, 1, 0>(ctx, P, tiles, smem, late_wait);
        case 6: return launch_q4k_t<6, 16, 1, 0>(ctx, P, tiles, smem, late_wait);
        case 7: return launch_q4k_t<7, 16, 1, 0>(ctx, P, tiles, smem, late_wait);
        case 8: return launch_q4k_t<8, 16, 1, 0>(ctx, P, tiles, smem, late_wait);
    }
    b200_set_error("gemv_q4k: nc=%d out of range", nc);
    return B200_E_ARG;
}
