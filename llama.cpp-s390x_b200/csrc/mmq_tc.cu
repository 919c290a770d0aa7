// mmq_tc.cu — prefill path (n > 8): tensor-core GEMM on tcgen05 with TMEM accumulators.
//
// Replaces, for large n, the reference's per-column vec_dot loop (ggml/src/ggml-cpu/ggml-cpu.c:1151-1439) and plays the
// role of ggml-cuda's mul_mat_q / cuBLAS branch (ggml/src/ggml-cuda/mmq.cuh:3446-3783, ggml-cuda.cu:1413-1530) — designed
// from the hardware up, not from those kernels:
//   * D[128 weight rows x NT activation columns] per CTA, accumulator in TMEM (fp32), one tcgen05.mma per 32 bytes of K;
//   * the weight tile (A operand) never exists in HBM in expanded form: 8 producer warps stream the raw quant blocks,
//     dequantise them in registers to fp16 (exact integer -> fp16 conversion by exponent-bias tricks, then ONE rounding
//     in `q * scale - min`) and store them straight into the 128-byte-swizzled K-major UMMA layout;
//   * the activation tile (B operand) is fetched by TMA from an fp16 copy of src1 that `tc_act_f16_kernel` writes once per
//     call (per-column power-of-two normalisation so neither overflow nor fp16 underflow can occur; undone exactly in
//     the epilogue).  Within every 4 consecutive K elements both operands use the order (0,2,1,3), which is what nibble
//     unpacking produces for free — a dot product does not care;
//   * one warp issues TMA, one warp issues MMA, 4-stage mbarrier ring; the producer warps double as epilogue warps
//     (tcgen05.ld -> scale -> coalesced f32 stores, or the MUL_MAT_ID scatter).
// Numerics: fp16 operands (2^-12 relative) + fp32 accumulation — closer to the exact product than the reference's own
// Q8_1 activation quantisation; NMSE vs ggml-cpu ~1e-5, far inside test-backend-ops' 5e-4.
#include <cuda.h>
#include <cuda_fp16.h>

#include "mm_problem.cuh"
#include "tc_ptx.cuh"

namespace {

constexpr int TC_M        = 128;     // weight rows per CTA (UMMA M)
constexpr int TC_KS       = 64;      // K elements per stage (one 128-byte swizzled row of fp16)
constexpr int TC_A_BYTES  = TC_M * 128;
constexpr int TC_PRODUCER = 256;     // 8 warps: thread -> (row = t & 127, half = t >> 7)
constexpr int TC_THREADS  = 512;     // 8 producer warps, TMA warp, MMA warp, 4 epilogue warps, 2 spare (whole warpgroups)

template <int NT> struct tc_cfg {
    static constexpr int STAGES  = NT == 256 ? 4 : 6;
    static constexpr int B_BYTES = NT * 128;
    static constexpr int SMEM    = STAGES * (TC_A_BYTES + B_BYTES) + 1024 /*alignment slack*/ + 256 /*barriers*/;
    static_assert(2 * NT <= 512, "two accumulators must fit the 512 TMEM columns");
};

struct tc_args {
    const uint8_t * W;
    int64_t nb01, nb02, nb03;
    int     m, k;
    int     n;                  // activation columns per batch
    int     ne12, r2, r3;       // batch dims (MM_PLAIN)
    float * Y;
    const float *   colscale;   // per activation position: 2^-e that undoes the fp16 normalisation
    const int32_t * ycol;       // per activation position: output column (MUL_MAT_ID scatter); NULL = identity
    int     mode;
    int     n_expert;
    const int32_t * expert_bounds;
    int64_t n_tiles;            // row tiles x column tiles (upper bound for MUL_MAT_ID) x batches
    int     dbg;                // timing experiments only: 1 no weight loads, 2 no convert/STS, 4 no TMA, 8 no MMA, 16 no stores
};

// ---- activations: f32 -> normalised fp16, K order (0,2,1,3) inside every group of 4 -------------------------------------
// one CTA per activation position p (plain: column i11 of batch; MUL_MAT_ID: sorted position -> pair -> (token, slot))
__global__ void __launch_bounds__(256) tc_act_f16_kernel(const mm_problem P, __half * __restrict__ Xh, float * __restrict__ colscale,
                                                         int32_t * __restrict__ ycol) {
    const int64_t p = blockIdx.x;
    const float * x;
    if (P.mode == MM_PLAIN) {
        const int64_t batch = p / P.ne11, i11 = p - batch * P.ne11;
        const int64_t i12 = batch % P.ne12, i13 = batch / P.ne12;
        x = reinterpret_cast<const float *>(reinterpret_cast<const char *>(P.X) + i11 * P.nb11 + i12 * P.nb12 + i13 * P.nb13);
    } else {
        const int     pair  = __ldg(P.ids_src1 + p);
        const int64_t token = pair / P.n_used;
        const int64_t i11   = (pair - token * P.n_used) % P.ne11;
        x = reinterpret_cast<const float *>(reinterpret_cast<const char *>(P.X) + i11 * P.nb11 + token * P.nb12);
        if (threadIdx.x == 0) ycol[p] = pair;
    }
    const int k4 = (int) (P.k / 4);
    float amax = 0.0f;
    for (int i = threadIdx.x; i < k4; i += 256) {
        const float4 v = __ldg(reinterpret_cast<const float4 *>(x) + i);
        amax = fmaxf(amax, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
    }
    __shared__ float red[8];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = amax;
    __syncthreads();
    amax = red[0];
#pragma unroll
    for (int w = 1; w < 8; ++w) amax = fmaxf(amax, red[w]);
    // amax = f * 2^e, f in [0.5, 1): scale by 2^(15 - e) so the column maximum lands in [2^14, 2^15)
    int e = 0;
    float up = 1.0f, down = 1.0f;
    if (amax > 0.0f && amax < INFINITY) {
        frexpf(amax, &e);
        int s = 15 - e;
        s = s > 126 ? 126 : (s < -126 ? -126 : s);
        up   = ldexpf(1.0f, s);
        down = ldexpf(1.0f, -s);
    }
    if (threadIdx.x == 0) colscale[p] = down;
    uint2 * out = reinterpret_cast<uint2 *>(Xh + p * P.k);
    for (int i = threadIdx.x; i < k4; i += 256) {
        const float4 v = __ldg(reinterpret_cast<const float4 *>(x) + i);
        const __half2 a = __floats2half2_rn(v.x * up, v.z * up);      // order (0, 2, 1, 3)
        const __half2 b = __floats2half2_rn(v.y * up, v.w * up);
        uint2 o;
        o.x = *reinterpret_cast<const uint32_t *>(&a);
        o.y = *reinterpret_cast<const uint32_t *>(&b);
        out[i] = o;
    }
}

// ---- weight fragments: what one producer thread holds for one stage (64 K elements of its row, its half of them) ----------
__device__ __forceinline__ uint32_t ld_nc_u32(const void * p) {
    uint32_t v; asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(v) : "l"(p)); return v;
}
__device__ __forceinline__ uint32_t ld_nc_u16(const void * p) {
    uint16_t v; asm volatile("ld.global.nc.u16 %0, [%1];" : "=h"(v) : "l"(p)); return v;
}
__device__ __forceinline__ uint4 ld_nc_u4(const void * p) {
    uint4 v; asm volatile("ld.global.nc.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p)); return v;
}
__device__ __forceinline__ uint32_t h2_bits(const __half2 h) { return *reinterpret_cast<const uint32_t *>(&h); }
__device__ __forceinline__ __half2  bits_h2(const uint32_t u) { return *reinterpret_cast<const __half2 *>(&u); }
__device__ __forceinline__ __half2  h2_const(const float v)  { return __float2half2_rn(v); }

// (b0, b2) and (b1, b3) low nibbles of a word as exact fp16 integers 1024 + q; the high nibbles as 1024 + 16 q
__device__ __forceinline__ void nib_pairs(const uint32_t w, __half2 & lo02, __half2 & lo13, __half2 & hi02, __half2 & hi13) {
    const uint32_t s = w >> 8;
    lo02 = bits_h2((w & 0x000F000Fu) | 0x64006400u);
    lo13 = bits_h2((s & 0x000F000Fu) | 0x64006400u);
    hi02 = bits_h2((w & 0x00F000F0u) | 0x64006400u);
    hi13 = bits_h2((s & 0x00F000F0u) | 0x64006400u);
}

template <int TYPE> struct w_frag;

// Q4_K (ggml/src/ggml-common.h:303-317, dequantize_row_q4_K ggml/src/ggml-quants.c:1467-1489): stage = 64-group g of
// super-block sb; thread half h owns qs bytes [16h, 16h+16) of the group: low nibbles -> sub-block 2g, high -> 2g+1.
template <> struct w_frag<B200_TYPE_Q4_K> {
    uint4 hdr, q;
    __device__ __forceinline__ void load(const uint8_t * row, const int it, const int h) {
        const uint8_t * sb = row + (int64_t) (it >> 2) * SZ_Q4_K;
        hdr = ld_nc_u4(sb);
        q   = ld_nc_u4(sb + 16 + 32 * (it & 3) + 16 * h);
    }
    // 4 chunks of 8 fp16; chunk slot indices within the 128-byte row returned in `slot`
    __device__ __forceinline__ void convert(const int it, const int h, uint4 (&out)[4], int (&slot)[4]) const {
        const int g = it & 3;
        int sc0, mn0, sc1, mn1;
        q4k_scale_min(2 * g,     hdr.y, hdr.z, hdr.w, sc0, mn0);
        q4k_scale_min(2 * g + 1, hdr.y, hdr.z, hdr.w, sc1, mn1);
        const float2 dm = __half22float2(bits_h2(hdr.x));
        const __half2 s_lo = h2_const(dm.x * (float) sc0), nb_lo = h2_const(-dm.y * (float) mn0);
        const __half2 s_hi = h2_const(dm.x * (float) sc1), nb_hi = h2_const(-dm.y * (float) mn1);
        const __half2 k1024 = h2_const(1024.0f), sixteenth = h2_const(0.0625f), m64 = h2_const(-64.0f);
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
        uint32_t lo[8], hi[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            __half2 a, b, c, d;
            nib_pairs(w[j], a, b, c, d);
            lo[2 * j]     = h2_bits(__hfma2(__hsub2(a, k1024), s_lo, nb_lo));
            lo[2 * j + 1] = h2_bits(__hfma2(__hsub2(b, k1024), s_lo, nb_lo));
            hi[2 * j]     = h2_bits(__hfma2(__hfma2(c, sixteenth, m64), s_hi, nb_hi));
            hi[2 * j + 1] = h2_bits(__hfma2(__hfma2(d, sixteenth, m64), s_hi, nb_hi));
        }
        // half 0 stores slots (0,1,4,5), half 1 stores (6,7,2,3): every 8-lane STS phase hits 8 distinct bank groups
        const uint4 l0 = make_uint4(lo[0], lo[1], lo[2], lo[3]), l1 = make_uint4(lo[4], lo[5], lo[6], lo[7]);
        const uint4 h0 = make_uint4(hi[0], hi[1], hi[2], hi[3]), h1 = make_uint4(hi[4], hi[5], hi[6], hi[7]);
        out[0] = h ? h0 : l0; slot[0] = h ? 6 : 0;
        out[1] = h ? h1 : l1; slot[1] = h ? 7 : 1;
        out[2] = h ? l0 : h0; slot[2] = h ? 2 : 4;
        out[3] = h ? l1 : h1; slot[3] = h ? 3 : 5;
    }
};

// Q4_0 (ggml-common.h:184-189, dequantize_row_q4_0 ggml-quants.c:314-332): stage = blocks 2it, 2it+1; thread half h owns
// block 2it+h: low nibbles = elements 0..15, high nibbles = elements 16..31, value (q - 8) * d.
template <> struct w_frag<B200_TYPE_Q4_0> {
    uint32_t w[5];
    __device__ __forceinline__ void load(const uint8_t * row, const int it, const int h) {
        const uint8_t * b = row + (int64_t) (2 * it + h) * SZ_Q4_0 - 2 * h;      // 4-byte aligned (18 * odd - 2)
#pragma unroll
        for (int j = 0; j < 4; ++j) w[j] = ld_nc_u32(b + 4 * j);
        w[4] = h ? ld_nc_u32(b + 16) : ld_nc_u16(b + 16);
    }
    __device__ __forceinline__ void convert(const int, const int h, uint4 (&out)[4], int (&slot)[4]) const {
        const __half2 d2 = __half2half2(__ushort_as_half((unsigned short) (h ? w[0] >> 16 : w[0] & 0xFFFF)));
        const __half2 k1032 = h2_const(1032.0f), sixteenth = h2_const(0.0625f), m72 = h2_const(-72.0f);
        uint32_t lo[8], hi[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t qs = h ? w[j + 1] : __funnelshift_r(w[j], w[j + 1], 16);
            __half2 a, b, c, d;
            nib_pairs(qs, a, b, c, d);
            lo[2 * j]     = h2_bits(__hmul2(__hsub2(a, k1032), d2));
            lo[2 * j + 1] = h2_bits(__hmul2(__hsub2(b, k1032), d2));
            hi[2 * j]     = h2_bits(__hmul2(__hfma2(c, sixteenth, m72), d2));
            hi[2 * j + 1] = h2_bits(__hmul2(__hfma2(d, sixteenth, m72), d2));
        }
        out[0] = make_uint4(lo[0], lo[1], lo[2], lo[3]); slot[0] = 4 * h;
        out[1] = make_uint4(lo[4], lo[5], lo[6], lo[7]); slot[1] = 4 * h + 1;
        out[2] = make_uint4(hi[0], hi[1], hi[2], hi[3]); slot[2] = 4 * h + 2;
        out[3] = make_uint4(hi[4], hi[5], hi[6], hi[7]); slot[3] = 4 * h + 3;
    }
};

// Q8_0 (ggml-common.h:241-246, dequantize_row_q8_0 ggml-quants.c:392-406): thread half h owns block 2it+h, value q * d.
template <> struct w_frag<B200_TYPE_Q8_0> {
    uint32_t w[9];
    __device__ __forceinline__ void load(const uint8_t * row, const int it, const int h) {
        const uint8_t * b = row + (int64_t) (2 * it + h) * SZ_Q8_0 - 2 * h;      // 4-byte aligned (34 * odd - 2)
#pragma unroll
        for (int j = 0; j < 8; ++j) w[j] = ld_nc_u32(b + 4 * j);
        w[8] = h ? ld_nc_u32(b + 32) : ld_nc_u16(b + 32);
    }
    __device__ __forceinline__ void convert(const int, const int h, uint4 (&out)[4], int (&slot)[4]) const {
        const __half2 d2 = __half2half2(__ushort_as_half((unsigned short) (h ? w[0] >> 16 : w[0] & 0xFFFF)));
        const __half2 k1152 = h2_const(1152.0f);
        uint32_t v[16];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t qs = (h ? w[j + 1] : __funnelshift_r(w[j], w[j + 1], 16)) ^ 0x80808080u;   // q + 128
            const __half2 a = bits_h2(__byte_perm(qs, 0x64646464u, 0x4240));   // (1024 + u0, 1024 + u2)
            const __half2 b = bits_h2(__byte_perm(qs, 0x64646464u, 0x4341));   // (1024 + u1, 1024 + u3)
            v[2 * j]     = h2_bits(__hmul2(__hsub2(a, k1152), d2));
            v[2 * j + 1] = h2_bits(__hmul2(__hsub2(b, k1152), d2));
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            out[c]  = make_uint4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
            slot[c] = 4 * h + c;
        }
    }
};


// ---- the GEMM ---------------------------------------------------------------------------------------------------------
// Persistent: CTA c works on tiles c, c + gridDim.x, ...  (row tile fastest, so CTAs running together share an activation
// tile in L2).  16 warps: 0-7 weight producers (4 pairs), 8 TMA, 9 MMA, 10-13 epilogue (one per TMEM lane quadrant).  Two
// TMEM accumulators, so the epilogue of tile i overlaps the main loop of tile i + 1.
struct tc_tile {
    const uint8_t * W;
    int row0, pos0, nvalid;
};

__device__ __forceinline__ bool tc_resolve(const tc_args & a, const int NT, int64_t T, tc_tile & t) {
    const int row_tiles = (a.m + TC_M - 1) / TC_M;
    const int rt = (int) (T % row_tiles);
    T /= row_tiles;
    t.row0 = rt * TC_M;
    t.W = a.W;
    if (a.mode == MM_PLAIN) {
        const int col_tiles = (a.n + NT - 1) / NT;
        const int ct = (int) (T % col_tiles), batch = (int) (T / col_tiles);
        const int i12 = batch % a.ne12, i13 = batch / a.ne12;
        t.W += (int64_t) (i12 / a.r2) * a.nb02 + (int64_t) (i13 / a.r3) * a.nb03;
        t.pos0   = batch * a.n + ct * NT;
        t.nvalid = min(NT, a.n - ct * NT);
        return true;
    }
    int ct = (int) T, e = 0, lo = 0, hi = 0;
    for (; e < a.n_expert; ++e) {
        lo = __ldg(a.expert_bounds + e);
        hi = __ldg(a.expert_bounds + e + 1);
        const int nt = (hi - lo + NT - 1) / NT;
        if (ct < nt) break;
        ct -= nt;
    }
    if (e >= a.n_expert) return false;
    t.W += (int64_t) e * a.nb02;
    t.pos0   = lo + ct * NT;
    t.nvalid = min(NT, hi - t.pos0);
    return true;
}

template <int TYPE, int NT>
__global__ void __launch_bounds__(TC_THREADS, 1) mmq_tc_f16_kernel(const __grid_constant__ CUtensorMap tmB, const tc_args a) {
    using cfg = tc_cfg<NT>;
    constexpr int STAGES = cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base  = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t sA    = base;
    const uint32_t sB    = base + STAGES * TC_A_BYTES;
    const uint32_t sBar  = sB + STAGES * cfg::B_BYTES;
    const uint32_t bFull = sBar, bEmpty = sBar + 8 * STAGES, bAccFull = sBar + 16 * STAGES, bAccEmpty = bAccFull + 16, sTmem = bAccEmpty + 16;
    uint8_t * gen_base   = smem_raw + (base - tc::smem_u32(smem_raw));

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nk = a.k / TC_KS;
    const int64_t n_tiles = a.n_tiles;

    if (warp == 8) {
        if (lane == 0) {
            tc::tma_prefetch_desc(&tmB);
            for (int s = 0; s < STAGES; ++s) {
                tc::mbar_init(bFull + 8 * s, 2 + 1);        // the 2 warps of a producer pair + the TMA thread's expect_tx
                tc::mbar_init(bEmpty + 8 * s, 1);           // one tcgen05.commit
            }
            for (int i = 0; i < 2; ++i) {
                tc::mbar_init(bAccFull + 8 * i, 1);         // tcgen05.commit after the last K step of a tile
                tc::mbar_init(bAccEmpty + 8 * i, 4);        // the 4 epilogue warps
            }
            tc::mbar_fence_init();
        }
        __syncwarp();
        tc::tmem_alloc(sTmem, 2 * NT);
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    uint32_t tmem;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(sTmem));

    if (warp < 8) {
        // ===== weight producers: raw blocks -> fp16 -> swizzled K-major A tile ==========================================
        // Warp pair p produces every 4th pipeline stage on its own (64 rows per warp, 4 (row, half) tasks per lane, adjacent
        // lanes share a row).  The cross-proxy fence drains a thread's outstanding loads, so the loads for the pair's next
        // stage are issued right after it and have a whole period (4 stage times) to land.
        const int pair = warp >> 1, h = lane & 1;
        const int rbase = 64 * (warp & 1) + (lane >> 1);
        uint32_t a_off[4];
        int rsw[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int r = rbase + 16 * j;
            a_off[j] = (uint32_t) ((r >> 3) * 1024 + (r & 7) * 128);
            rsw[j]   = r & 7;
        }
        w_frag<TYPE> f[4];
        // g = running stage index over all tiles of this CTA; (tile ordinal, k step) = (g / nk, g % nk)
        int64_t T = blockIdx.x;
        tc_tile t;
        while (T < n_tiles && !tc_resolve(a, NT, T, t)) T += gridDim.x;
        int it = pair;                              // k step of the fragment held in f (tile T)
        while (T < n_tiles && it >= nk) {           // nk < 4: this pair may have nothing in a tile
            it -= nk;
            do { T += gridDim.x; } while (T < n_tiles && !tc_resolve(a, NT, T, t));
        }
        int64_t g = 0;                              // stages of earlier tiles
        int64_t gcur = pair;                        // global stage index of f
        (void) g;
        if (T < n_tiles && !(a.dbg & 1)) {
#pragma unroll
            for (int j = 0; j < 4; ++j) f[j].load(t.W + (int64_t) min(t.row0 + rbase + 16 * j, a.m - 1) * a.nb01, it, h);
        }
        while (T < n_tiles) {
            const int s = (int) (gcur % STAGES);
            tc::mbar_wait(bEmpty + 8 * s, (uint32_t) ((gcur / STAGES) & 1) ^ 1);
            uint8_t * dst = gen_base + s * TC_A_BYTES;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (a.dbg & 2) break;
                uint4 out[4]; int slot[4];
                f[j].convert(it, h, out, slot);
#pragma unroll
                for (int c = 0; c < 4; ++c) *reinterpret_cast<uint4 *>(dst + a_off[j] + ((slot[c] ^ rsw[j]) << 4)) = out[c];
            }
            tc::fence_proxy_async();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(bFull + 8 * s);
            // advance to this pair's next stage (possibly in a later tile) and start its loads
            it += 4; gcur += 4;
            while (T < n_tiles && it >= nk) {
                it -= nk;
                do { T += gridDim.x; } while (T < n_tiles && !tc_resolve(a, NT, T, t));
            }
            if (T < n_tiles && !(a.dbg & 1)) {
#pragma unroll
                for (int j = 0; j < 4; ++j) f[j].load(t.W + (int64_t) min(t.row0 + rbase + 16 * j, a.m - 1) * a.nb01, it, h);
            }
        }
    } else if (warp == 8) {
        // ===== TMA: activation tiles ========================================================================================
        if (lane == 0) {
            int64_t gcur = 0;
            tc_tile t;
            for (int64_t T = blockIdx.x; T < n_tiles; T += gridDim.x) {
                if (!tc_resolve(a, NT, T, t)) continue;
                for (int it = 0; it < nk; ++it, ++gcur) {
                    const int s = (int) (gcur % STAGES);
                    tc::mbar_wait(bEmpty + 8 * s, (uint32_t) ((gcur / STAGES) & 1) ^ 1);
                    if (a.dbg & 4) { tc::mbar_arrive(bFull + 8 * s); continue; }
                    tc::mbar_arrive_expect_tx(bFull + 8 * s, cfg::B_BYTES);
                    tc::tma_load_2d(sB + s * cfg::B_BYTES, &tmB, bFull + 8 * s, it * TC_KS, t.pos0);
                }
            }
        }
    } else if (warp == 9) {
        // ===== MMA issue ====================================================================================================
        if (lane == 0) {
            constexpr uint32_t idesc = tc::umma_idesc(/*f32*/ 1, /*f16*/ 0, /*f16*/ 0, TC_M, NT);
            int64_t gcur = 0;
            int     ti = 0;                          // ordinal of the tile within this CTA
            tc_tile t;
            for (int64_t T = blockIdx.x; T < n_tiles; T += gridDim.x) {
                if (!tc_resolve(a, NT, T, t)) continue;
                const int acc = ti & 1;
                tc::mbar_wait(bAccEmpty + 8 * acc, (uint32_t) ((ti >> 1) & 1) ^ 1);
                tc::tc_fence_after();
                for (int it = 0; it < nk; ++it, ++gcur) {
                    const int s = (int) (gcur % STAGES);
                    tc::mbar_wait(bFull + 8 * s, (uint32_t) ((gcur / STAGES) & 1));
                    tc::tc_fence_after();
                    const uint64_t da = tc::umma_desc_sw128(sA + s * TC_A_BYTES);
                    const uint64_t db = tc::umma_desc_sw128(sB + s * cfg::B_BYTES);
#pragma unroll
                    for (int kk = 0; kk < TC_KS / 16; ++kk) {
                        if (a.dbg & 8) break;
                        tc::umma_f16(tmem + acc * NT, da + 2 * kk, db + 2 * kk, idesc, (it | kk) != 0);
                    }
                    tc::tc_commit(bEmpty + 8 * s);
                }
                tc::tc_commit(bAccFull + 8 * acc);
                ++ti;
            }
        }
    } else if (warp < 14) {
        // ===== epilogue: TMEM -> registers -> global (coalesced along rows; MUL_MAT_ID scatters columns) ===================
        const int q = warp & 3;
        int ti = 0;
        tc_tile t;
        for (int64_t T = blockIdx.x; T < n_tiles; T += gridDim.x) {
            if (!tc_resolve(a, NT, T, t)) continue;
            const int acc = ti & 1;
            tc::mbar_wait(bAccFull + 8 * acc, (uint32_t) ((ti >> 1) & 1));
            tc::tc_fence_after();
            const int row = t.row0 + 32 * q + lane;
#pragma unroll 1
            for (int c0 = 0; c0 < NT; c0 += 32) {
                if (c0 >= t.nvalid) break;                            // warp-uniform
                uint32_t v[32];
                tc::tmem_ld32(tmem + ((uint32_t) (32 * q) << 16) + (uint32_t) (acc * NT + c0), v);
                // this lane fetches the scale / destination column of position c0 + lane; shuffles hand them round
                const int  pl  = t.pos0 + c0 + lane;
                const bool pok = c0 + lane < t.nvalid;
                const float   my_sc = pok ? __ldg(a.colscale + pl) : 0.0f;
                const int32_t my_yc = pok ? (a.ycol ? __ldg(a.ycol + pl) : pl) : 0;
                tc::tmem_ld_wait();
                if (!(a.dbg & 16)) {
#pragma unroll
                    for (int c = 0; c < 32; ++c) {
                        const float   sc = __shfl_sync(0xffffffffu, my_sc, c);
                        const int32_t yc = __shfl_sync(0xffffffffu, my_yc, c);
                        if (c0 + c < t.nvalid && row < a.m) a.Y[(int64_t) yc * a.m + row] = __uint_as_float(v[c]) * sc;
                    }
                }
            }
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(bAccEmpty + 8 * acc);
            ++ti;
        }
    }
    __syncthreads();
    if (warp == 8) {
        __syncwarp();
        tc::tc_fence_after();
        tc::tmem_dealloc(tmem, 2 * NT);
    }
}

// ---- host ---------------------------------------------------------------------------------------------------------------
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

encode_tiled_fn get_encode() {
    static encode_tiled_fn fn = [] {
        void * p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) p = nullptr;
        return (encode_tiled_fn) p;
    }();
    return fn;
}

bool tc_type_ok(int type) { return type == B200_TYPE_Q4_K || type == B200_TYPE_Q4_0 || type == B200_TYPE_Q8_0; }

template <int TYPE, int NT>
int launch_one(b200_ctx * ctx, const CUtensorMap & tm, const tc_args & a) {
    static bool configured[16] = {};
    auto kern = mmq_tc_f16_kernel<TYPE, NT>;
    if (!configured[ctx->device & 15]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, tc_cfg<NT>::SMEM));
        configured[ctx->device & 15] = true;
    }
    const unsigned grid = (unsigned) (a.n_tiles < ctx->sm_count ? a.n_tiles : ctx->sm_count);
    kern<<<grid, TC_THREADS, tc_cfg<NT>::SMEM, ctx->stream>>>(tm, a);
    B200_CUDA_TRY(cudaGetLastError());
    ctx->launches++;
    return B200_OK;
}

template <int NT>
int launch_type(b200_ctx * ctx, int type, const CUtensorMap & tm, const tc_args & a) {
    switch (type) {
        case B200_TYPE_Q4_K: return launch_one<B200_TYPE_Q4_K, NT>(ctx, tm, a);
        case B200_TYPE_Q4_0: return launch_one<B200_TYPE_Q4_0, NT>(ctx, tm, a);
        case B200_TYPE_Q8_0: return launch_one<B200_TYPE_Q8_0, NT>(ctx, tm, a);
    }
    b200_set_error("tcgen05 GEMM: unsupported weight type %d", type);
    return B200_E_UNSUPPORTED;
}

}  // namespace

int b200_tc_supported(const b200_ctx * ctx, int type, const mm_problem & P) {
    if (!tc_type_ok(type) || !get_encode()) return 0;
    if (P.k % TC_KS != 0 || P.k > (1 << 24) || P.m > (1 << 30)) return 0;
    const uintptr_t bits = (uintptr_t) P.W | (uintptr_t) P.nb01 | (uintptr_t) P.nb02 | (uintptr_t) P.nb03;
    if (bits & (type == B200_TYPE_Q4_K ? 15 : 3)) return 0;
    if (((uintptr_t) P.X | (uintptr_t) P.nb11 | (uintptr_t) P.nb12 | (uintptr_t) P.nb13) & 15) return 0;
    if (P.mode == MM_ID_PAIRS) return 0;
    if (tc_cfg<256>::SMEM > ctx->smem_optin) return 0;
    return 1;
}

// n_pos = number of activation positions (MM_PLAIN: ne11 * batches; MM_ID_SORTED: pairs).  Uses the context scratch
// AFTER `ws_used` bytes (the caller may keep ids there).
int b200_launch_gemm_tc(b200_ctx * ctx, int type, const mm_problem & P, int64_t n_pos, size_t ws_used) {
    const size_t xh_bytes = (size_t) n_pos * P.k * sizeof(__half);
    const size_t off_xh   = (ws_used + 1023) / 1024 * 1024;
    const size_t off_sc   = off_xh + (xh_bytes + 255) / 256 * 256;
    const size_t off_yc   = off_sc + ((size_t) n_pos * 4 + 255) / 256 * 256;
    const size_t total    = off_yc + ((size_t) n_pos * 4 + 255) / 256 * 256;
    if ((size_t) ctx->ws_size < total) { b200_set_error("tcgen05 GEMM: scratch not reserved"); return B200_E_ARG; }
    uint8_t * ws = (uint8_t *) ctx->ws;
    __half *  Xh = (__half *) (ws + off_xh);
    float *   sc = (float *) (ws + off_sc);
    int32_t * yc = (int32_t *) (ws + off_yc);

    tc_act_f16_kernel<<<(unsigned) n_pos, 256, 0, ctx->stream>>>(P, Xh, sc, yc);
    B200_CUDA_TRY(cudaGetLastError());
    ctx->launches++;

    const int64_t batches = P.mode == MM_PLAIN ? P.ne12 * P.ne13 : 1;
    const int64_t n       = P.mode == MM_PLAIN ? P.ne11 : n_pos;
    const int     NT      = n <= 128 ? 128 : 256;

    CUtensorMap tm;
    const cuuint64_t gdim[2]    = {(cuuint64_t) P.k, (cuuint64_t) n_pos};
    const cuuint64_t gstride[1] = {(cuuint64_t) P.k * sizeof(__half)};
    const int64_t row_tiles = (P.m + TC_M - 1) / TC_M;
    const cuuint32_t box[2]     = {(cuuint32_t) TC_KS, (cuuint32_t) NT};
    const cuuint32_t estr[2]    = {1, 1};
    const CUresult cr = get_encode()(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, Xh, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) { b200_set_error("cuTensorMapEncodeTiled failed (%d)", (int) cr); return B200_E_CUDA; }

    tc_args a = {};
    a.W = P.W; a.nb01 = P.nb01; a.nb02 = P.nb02; a.nb03 = P.nb03;
    a.m = (int) P.m; a.k = (int) P.k; a.n = (int) n;
    a.ne12 = (int) P.ne12; a.r2 = (int) P.r2; a.r3 = (int) P.r3;
    a.Y = P.Y; a.colscale = sc; a.ycol = P.mode == MM_PLAIN ? nullptr : yc;
    a.mode = P.mode; a.n_expert = P.n_expert; a.expert_bounds = P.expert_bounds;
    a.dbg = ctx->tc_dbg;

    const int64_t col_tiles = P.mode == MM_PLAIN ? (n + NT - 1) / NT : (n_pos + NT - 1) / NT + P.n_expert;
    a.n_tiles = row_tiles * col_tiles * batches;
    return NT == 128 ? launch_type<128>(ctx, type, tm, a) : launch_type<256>(ctx, type, tm, a);
}

size_t b200_gemm_tc_scratch(const mm_problem & P, int64_t n_pos, size_t ws_used) {
    const size_t xh_bytes = (size_t) n_pos * P.k * sizeof(__half);
    const size_t off_xh   = (ws_used + 1023) / 1024 * 1024;
    return off_xh + (xh_bytes + 255) / 256 * 256 + 2 * (((size_t) n_pos * 4 + 255) / 256 * 256);
}
