// mmq_tc.cu — prefill path (n > 8): tensor-core GEMM on tcgen05 with the accumulators AND the weight operand in tensor memory.
//
// Replaces, for large n, the reference's per-column vec_dot loop (ggml/src/ggml-cpu/ggml-cpu.c:1151-1439) and plays the
// role of ggml-cuda's mul_mat_q / cuBLAS branch (ggml/src/ggml-cuda/mmq.cuh:3446-3783, ggml-cuda.cu:1413-1530) — designed
// from the hardware up, not from those kernels:
//   * persistent CTAs; D[128 weight rows x up to 192 activation columns] per tile, fp32 accumulator in TMEM, two accumulators
//     so the epilogue of one tile overlaps the main loop of the next;
//   * the weight tile (A operand) never exists in HBM in expanded form, and not in shared memory either: TMA streams the RAW
//     quant blocks (a box of 128 rows x 144 bytes) into a shared-memory ring, 8 converter warps dequantise them to fp16 (exact
//     integer -> fp16 conversion by exponent-bias tricks, then ONE rounding in `q * scale - min`) and write them with
//     tcgen05.st into a 4-stage ring of TENSOR MEMORY columns (thread = tile row = TMEM lane, 32 columns = 64 K elements); the
//     MMA takes A from there (`tcgen05.mma [d], [a_tmem], b_desc`).  Round 1 kept A in shared memory: 16 KB written + 16 KB read
//     per 64-K stage out of ~107 KB of shared-memory traffic, which is what bounded that kernel (836 clocks per stage against
//     512 of MMA time); tools/tc_ts_probe.cu checks the TMEM operand layout and shows the MMA keeps its full rate from it;
//   * the activation tile (B operand) is fetched by TMA (boxes of 64 columns x 64 K) from an fp16 copy of src1 that
//     `tc_act_f16_kernel` writes once per call (per-column power-of-two normalisation so neither overflow nor fp16 underflow
//     can occur; undone exactly in the epilogue).  Within every 4 consecutive K elements both operands use the order
//     (0,2,1,3), which is what nibble unpacking produces for free — a dot product does not care;
//   * tiles are as wide as their valid columns: the MMA's N is the tile's column count rounded up to 16 and only the boxes that
//     hold valid columns are fetched, so a ragged last tile or a MUL_MAT_ID expert with few tokens costs what it uses;
//   * warp roles: 0-11 converters (three groups of four, a group takes every third stage), 12 activation TMA, 13 MMA issue, 14-17
//     epilogue (tcgen05.ld -> scale -> coalesced f32 stores, or the MUL_MAT_ID scatter), 18 weight TMA (its own warp: the raw ring
//     holds 16 stages and must run that far ahead — HBM latency — while the activation ring holds 6; one warp serving both in
//     program order let the raw blocks arrive late, ncu: 26 % of the converters' time in the raw-slot wait); mbarrier rings:
//     raw weights, A stages (TMEM), B tiles, accumulators.  The converters are what the stage rate hangs on (ncu source page, profiles/): ~290 instructions
//     per warp and stage in a mostly serial chain, so the third group is what keeps the A ring ahead of the MMA warp.
// Numerics: fp16 operands (2^-12 relative) + fp32 accumulation — closer to the exact product than the reference's own
// Q8_1 activation quantisation; NMSE vs ggml-cpu ~1.5e-5 (that IS the Q8_1 noise), far inside test-backend-ops' 5e-4.  The sum
// over K of one output element is formed in the same order whatever the tile width, so results do not depend on the tiling.
#include <cuda.h>
#include <cuda_fp16.h>
#include <atomic>
#include <cstdio>

#include "mm_problem.cuh"
#include "tc_ptx.cuh"

namespace {

constexpr int TC_M        = 128;     // weight rows per CTA tile (UMMA M) = TMEM lanes
constexpr int TC_KS       = 64;      // K elements per MMA stage (one 128-byte swizzled row of fp16 B; 32 TMEM columns of A)
constexpr int TC_NT_MAX   = 192;     // widest tile: 2 accumulators x 192 columns + the A ring = the 512 TMEM columns
constexpr int TC_A_SLOTS  = 4;       // A stages in tensor memory
constexpr int TC_A_COL0   = 2 * TC_NT_MAX;
constexpr int TC_A_COLS   = TC_KS / 2;
constexpr int TC_B_BOX    = 64;      // activation columns per TMA box (CTA pair: 32 — a CTA holds half of the tile's columns)
constexpr int TC_B_BOX_BYTES = TC_B_BOX * 128;
constexpr int TC_B_BOX_P  = 32, TC_B_BOX_P_BYTES = TC_B_BOX_P * 128;
constexpr int TC_B_BYTES  = TC_NT_MAX * 128;
constexpr int TC_RAW_ROW  = 144;     // bytes of raw blocks per row and raw slot (Q4_K: 1 super-block, Q4_0: 8 blocks, Q8_0: 4 + pad)
constexpr int TC_CONV_GROUPS = 3;    // converter groups of 4 warps (one warp per TMEM lane quadrant); group g takes every 3rd stage
constexpr int TC_CONV_WARPS  = 4 * TC_CONV_GROUPS;
constexpr int TC_W_TMA = TC_CONV_WARPS, TC_W_MMA = TC_CONV_WARPS + 1, TC_W_EPI = TC_CONV_WARPS + 2, TC_W_RAW = TC_CONV_WARPS + 6;   // warp roles
constexpr int TC_THREADS  = 32 * (TC_CONV_WARPS + 7);   // 12 converter warps, activation-TMA warp, MMA warp, 4 epilogue warps, weight-TMA warp
static_assert(TC_A_COL0 + TC_A_SLOTS * TC_A_COLS <= 512, "accumulators and the A ring must fit the 512 TMEM columns");
static_assert((TC_A_SLOTS & (TC_A_SLOTS - 1)) == 0, "A ring slots are taken from the stage counter by masking");

// RAW_ROW: bytes per row and raw slot — 144 (Q4_K: 1 super-block, Q4_0: 8 blocks, Q8_0: 4 blocks + pad) or 224 (Q6_K: one 210-byte
// super-block + pad; the wider raw ring is paid for with one activation slot and one raw slot)
template <int RAW_ROW = TC_RAW_ROW> struct tc_cfg {
    static constexpr int B_SLOTS   = RAW_ROW > TC_RAW_ROW ? 4 : 6;           // an even number: stages travel in pairs
    static constexpr int RAW_SLOTS = RAW_ROW > TC_RAW_ROW ? 3 : 4;
    static constexpr int RAW_BYTES = TC_M * RAW_ROW;
    static constexpr int SMEM      = B_SLOTS * TC_B_BYTES + RAW_SLOTS * RAW_BYTES + 1024 /*alignment slack*/ + 512 /*barriers*/;
    static_assert(SMEM <= 232448, "exceeds the 227 KB a CTA can opt into");
};

// K elements covered by one raw slot (144 bytes per row)
template <int TYPE> struct raw_k          { static constexpr int value = 256; };
template <> struct raw_k<B200_TYPE_Q8_0>  { static constexpr int value = 128; };
template <> struct raw_k<B200_TYPE_F16>   { static constexpr int value = 64; };        // fp16 "weights" (see w_frag<F16>): one stage per slot
// bytes by which raw slot i of a row is fetched early so that the TMA source address is 16-byte aligned
template <int TYPE> __device__ __forceinline__ uint32_t raw_skew(const int i) {
    return TYPE == B200_TYPE_Q8_0 ? 8u * (uint32_t) (i & 1) : (TYPE == B200_TYPE_Q6_K ? 2u * (uint32_t) (i & 7) : 0u);
}
template <int TYPE> struct raw_row        { static constexpr int value = TC_RAW_ROW; };
template <> struct raw_row<B200_TYPE_Q6_K> { static constexpr int value = 224; };
template <> struct raw_row<B200_TYPE_Q5_K> { static constexpr int value = SZ_Q5_K; };       // 176: a whole number of 16-byte units, no skew
template <> struct raw_row<B200_TYPE_F16>  { static constexpr int value = 2 * TC_KS; };    // 128: the 64 halves of a stage
constexpr int raw_row_of(const int type) { return type == B200_TYPE_Q6_K ? 224 : (type == B200_TYPE_Q5_K ? SZ_Q5_K : (type == B200_TYPE_F16 ? 2 * TC_KS : TC_RAW_ROW)); }
constexpr int raw_k_of(const int type)   { return type == B200_TYPE_Q8_0 ? 128 : (type == B200_TYPE_F16 ? 64 : 256); }

struct tc_args {
    const uint8_t * W;
    int64_t nb01, nb02, nb03;
    int     m, k;
    int     n;                  // activation columns per batch
    int     nt;                 // tile width in columns (64, 128 or 192); the last tile of a row of tiles may be narrower
    int     col_tiles;          // MM_PLAIN: tiles per batch along n
    int     ne12, r2, r3;       // batch dims (MM_PLAIN)
    int     ne02;               // weight matrices per ne03 slice
    int     raw_step;           // 16-bit elements of a weight row per raw slot (row bytes of raw_k K elements / 2)
    uint32_t k6400;             // 0x64006400 (see and_or)
    float * Y;
    const float *   colscale;   // per activation position: 2^-e that undoes the fp16 normalisation
    const int32_t * ycol;       // per activation position: output column (MUL_MAT_ID scatter); NULL = identity
    int     mode;
    int     n_expert;
    const int32_t * expert_bounds;
    // several weight matrices against the same activations (q/k/v, gate/up; MM_PLAIN, no batch dims): row tiles are numbered
    // through all of them, every matrix padded to whole tiles; matrix j has its own TMA map, height and output
    int     n_mats;
    int     rt_end[3];          // cumulative row tiles
    int     m_j[3];
    float * Y_j[3];
    int64_t n_tiles;            // row tiles x column tiles (upper bound for MUL_MAT_ID) x batches
    int     grid;               // persistent CTAs
    int     split_k;            // K parts per tile (1, 2 or 4); the parts of a tile are summed in ascending K order
    int32_t * tile_sync;        //    through these per-(tile, lane quadrant) turnstiles (zeroed per launch)
    // row split: the all-gather of the output rows is fused into the epilogue — every result also goes to
    // peer_Y[p][col * peer_m + peer_row0 + row] of each GPU (peer-mapped pointers, own GPU included if listed)
    int     n_peers;
    float * peer_Y[8];
    int64_t peer_m, peer_row0;
};

// ---- activations: f32 -> normalised fp16, K order (0,2,1,3) inside every group of 4 -------------------------------------
// one CTA per activation position p (plain: column i11 of batch; MUL_MAT_ID: sorted position -> pair -> (token, slot))
__global__ void __launch_bounds__(256) tc_act_f16_kernel(const mm_problem P, __half * __restrict__ Xh, float * __restrict__ colscale,
                                                         int32_t * __restrict__ ycol) {
    // programmatic dependent launch (ctx option tc_pdl): this grid may start while the previous kernel of the stream (normally the
    // previous GEMM, which still reads the fp16 scratch this kernel rewrites, and may be producing X) is finishing — wait for it
    // here; the GEMM that consumes this kernel's output may be launched right away (it sets up barriers / TMEM, then waits for
    // this grid).  Both instructions are no-ops in a launch without the attribute.
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
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
    // the first 8 float4 of every thread (all of a column up to k = 8192) stay in registers between the two passes: all loads
    // of a thread are in flight together and the second pass does not go back to L2
    constexpr int KEEP = 8;
    float4 keep[KEEP];
#pragma unroll
    for (int j = 0; j < KEEP; ++j) {
        const int i = threadIdx.x + 256 * j;
        keep[j] = i < k4 ? __ldg(reinterpret_cast<const float4 *>(x) + i) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
#pragma unroll
    for (int j = 0; j < KEEP; ++j) amax = fmaxf(amax, fmaxf(fmaxf(fabsf(keep[j].x), fabsf(keep[j].y)), fmaxf(fabsf(keep[j].z), fabsf(keep[j].w))));
    for (int i = threadIdx.x + 256 * KEEP; i < k4; i += 256) {
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
    auto emit = [&](const int i, const float4 v) {
        const __half2 a = __floats2half2_rn(v.x * up, v.z * up);      // order (0, 2, 1, 3)
        const __half2 b = __floats2half2_rn(v.y * up, v.w * up);
        uint2 o;
        o.x = *reinterpret_cast<const uint32_t *>(&a);
        o.y = *reinterpret_cast<const uint32_t *>(&b);
        out[i] = o;
    };
#pragma unroll
    for (int j = 0; j < KEEP; ++j) {
        const int i = threadIdx.x + 256 * j;
        if (i < k4) emit(i, keep[j]);
    }
    for (int i = threadIdx.x + 256 * KEEP; i < k4; i += 256) emit(i, __ldg(reinterpret_cast<const float4 *>(x) + i));
}

// the same for short columns (k <= 1024: the q rows and the score rows of prompt-time attention, tens of thousands of positions): one warp
// per position, eight positions per CTA, the column in registers, no block-wide synchronisation
__global__ void __launch_bounds__(256) tc_act_f16_warp_kernel(const mm_problem P, __half * __restrict__ Xh, float * __restrict__ colscale,
                                                              int32_t * __restrict__ ycol, const int64_t n_pos) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int lane = threadIdx.x & 31;
    const int64_t p = (int64_t) blockIdx.x * 8 + (threadIdx.x >> 5);
    if (p >= n_pos) return;
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
        if (lane == 0) ycol[p] = pair;
    }
    const int k4 = (int) (P.k / 4);                       // <= 256: at most 8 float4 per lane
    float4 keep[8];
    float amax = 0.0f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int i = lane + 32 * j;
        keep[j] = i < k4 ? __ldg(reinterpret_cast<const float4 *>(x) + i) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        amax = fmaxf(amax, fmaxf(fmaxf(fabsf(keep[j].x), fabsf(keep[j].y)), fmaxf(fabsf(keep[j].z), fabsf(keep[j].w))));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    int e = 0;
    float up = 1.0f, down = 1.0f;
    if (amax > 0.0f && amax < INFINITY) {                 // (as in tc_act_f16_kernel)
        frexpf(amax, &e);
        int s = 15 - e;
        s = s > 126 ? 126 : (s < -126 ? -126 : s);
        up   = ldexpf(1.0f, s);
        down = ldexpf(1.0f, -s);
    }
    if (lane == 0) colscale[p] = down;
    uint2 * out = reinterpret_cast<uint2 *>(Xh + p * P.k);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int i = lane + 32 * j;
        if (i < k4) {
            const __half2 a = __floats2half2_rn(keep[j].x * up, keep[j].z * up);      // order (0, 2, 1, 3)
            const __half2 b = __floats2half2_rn(keep[j].y * up, keep[j].w * up);
            uint2 o;
            o.x = *reinterpret_cast<const uint32_t *>(&a);
            o.y = *reinterpret_cast<const uint32_t *>(&b);
            out[i] = o;
        }
    }
}

// ---- weight fragments: what one converter thread handles for one stage (64 K elements of its row, its half of them) ------
__device__ __forceinline__ uint32_t lds_u32(const uint32_t a) {
    uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v;
}
__device__ __forceinline__ uint4 lds_u4(const uint32_t a) {
    uint4 v; asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a)); return v;
}
__device__ __forceinline__ uint32_t h2_bits(const __half2 h) { return *reinterpret_cast<const uint32_t *>(&h); }
__device__ __forceinline__ __half2  bits_h2(const uint32_t u) { return *reinterpret_cast<const __half2 *>(&u); }
__device__ __forceinline__ __half2  h2_const(const float v)  { return __float2half2_rn(v); }

// (a & mask) | k in ONE instruction: LOP3 takes a single immediate, so the exponent pattern 0x64006400 has to sit in a register — it
// is read from the kernel arguments (tc_args::k6400), which keeps ptxas from folding it back into a second immediate and splitting
// the operation in two (the ALU pipe is what the converters hang on: ncu, 59 % busy at 0.5 instructions / clock)
__device__ __forceinline__ uint32_t and_or(const uint32_t a, const uint32_t mask, const uint32_t k) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(a), "r"(mask), "r"(k));
    return r;
}
// (b0, b2) and (b1, b3) low nibbles of a word as exact fp16 integers 1024 + q; the high nibbles as 1024 + 16 q
__device__ __forceinline__ void nib_pairs(const uint32_t w, const uint32_t k6400, __half2 & lo02, __half2 & lo13, __half2 & hi02, __half2 & hi13) {
    const uint32_t s = w >> 8;
    lo02 = bits_h2(and_or(w, 0x000F000Fu, k6400));
    lo13 = bits_h2(and_or(s, 0x000F000Fu, k6400));
    hi02 = bits_h2(and_or(w, 0x00F000F0u, k6400));
    hi13 = bits_h2(and_or(s, 0x00F000F0u, k6400));
}

template <int TYPE> struct w_frag;

// Q4_K (ggml/src/ggml-common.h:303-317, dequantize_row_q4_K ggml/src/ggml-quants.c:1467-1489): stage = 64-group g of
// super-block sb; thread half h owns qs bytes [16h, 16h+16) of the group: low nibbles -> sub-block 2g, high -> 2g+1.
template <> struct w_frag<B200_TYPE_Q4_K> {
    uint4 hdr, q;
    // row = shared-memory address of this row's super-block; sub = 64-group inside it
    __device__ __forceinline__ void load(const uint32_t row, const int sub, const int h) {
        hdr = lds_u4(row);
        q   = lds_u4(row + 16 + 32 * sub + 16 * h);
    }
    // both halves of a row's stage: one header load (the scale / min work of the two halves is then common code)
    static __device__ __forceinline__ void load_row(w_frag & f0, w_frag & f1, const uint32_t row, const int sub) {
        f0.hdr = lds_u4(row); f1.hdr = f0.hdr;
        f0.q = lds_u4(row + 16 + 32 * sub); f1.q = lds_u4(row + 16 + 32 * sub + 16);
    }
    // 4 chunks of 8 fp16; chunk slot indices within the 128-byte row returned in `slot`
    __device__ __forceinline__ void convert(const int sub, const int h, const uint32_t k6400, uint4 (&out)[4], int (&slot)[4]) const {
        const int g = sub;
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
            nib_pairs(w[j], k6400, a, b, c, d);
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

// Q5_K (ggml-common.h:330-346, dequantize_row_q5_K ggml-quants.c:1669-1697): Q4_K's layout with 32 bytes qh between the scales and the
// nibbles — bit j of qh[l] is the fifth bit of element l of sub-block j.  Same stage / half / slot assignment as Q4_K.
template <> struct w_frag<B200_TYPE_Q5_K> {
    uint4 hdr, q, qh;
    __device__ __forceinline__ void load(const uint32_t row, const int sub, const int h) {
        hdr = lds_u4(row);
        qh  = lds_u4(row + 16 + 16 * h);
        q   = lds_u4(row + 48 + 32 * sub + 16 * h);
    }
    static __device__ __forceinline__ void load_row(w_frag & f0, w_frag & f1, const uint32_t row, const int sub) {
        f0.hdr = lds_u4(row); f1.hdr = f0.hdr;
        f0.qh = lds_u4(row + 16); f1.qh = lds_u4(row + 32);
        f0.q = lds_u4(row + 48 + 32 * sub); f1.q = lds_u4(row + 48 + 32 * sub + 16);
    }
    __device__ __forceinline__ void convert(const int sub, const int h, const uint32_t k6400, uint4 (&out)[4], int (&slot)[4]) const {
        (void) k6400;
        const int g = sub;
        int sc0, mn0, sc1, mn1;
        q4k_scale_min(2 * g,     hdr.y, hdr.z, hdr.w, sc0, mn0);
        q4k_scale_min(2 * g + 1, hdr.y, hdr.z, hdr.w, sc1, mn1);
        const float2 dm = __half22float2(bits_h2(hdr.x));
        const __half2 s_lo = h2_const(dm.x * (float) sc0), nb_lo = h2_const(-dm.y * (float) mn0);
        const __half2 s_hi = h2_const(dm.x * (float) sc1), nb_hi = h2_const(-dm.y * (float) mn1);
        const __half2 k1024 = h2_const(1024.0f);
        const uint32_t w[4] = {q.x, q.y, q.z, q.w}, hw[4] = {qh.x, qh.y, qh.z, qh.w};
        uint32_t lo[8], hi[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t ql = (w[j] & 0x0F0F0F0Fu)        | (((hw[j] >> (2 * g)) & 0x01010101u) << 4);          // 4 x (0..31)
            const uint32_t qu = ((w[j] >> 4) & 0x0F0F0F0Fu) | (((hw[j] >> (2 * g + 1)) & 0x01010101u) << 4);
            lo[2 * j]     = h2_bits(__hfma2(__hsub2(bits_h2(__byte_perm(ql, 0x64646464u, 0x4240)), k1024), s_lo, nb_lo));
            lo[2 * j + 1] = h2_bits(__hfma2(__hsub2(bits_h2(__byte_perm(ql, 0x64646464u, 0x4341)), k1024), s_lo, nb_lo));
            hi[2 * j]     = h2_bits(__hfma2(__hsub2(bits_h2(__byte_perm(qu, 0x64646464u, 0x4240)), k1024), s_hi, nb_hi));
            hi[2 * j + 1] = h2_bits(__hfma2(__hsub2(bits_h2(__byte_perm(qu, 0x64646464u, 0x4341)), k1024), s_hi, nb_hi));
        }
        const uint4 l0 = make_uint4(lo[0], lo[1], lo[2], lo[3]), l1 = make_uint4(lo[4], lo[5], lo[6], lo[7]);
        const uint4 h0 = make_uint4(hi[0], hi[1], hi[2], hi[3]), h1 = make_uint4(hi[4], hi[5], hi[6], hi[7]);
        out[0] = h ? h0 : l0; slot[0] = h ? 6 : 0;
        out[1] = h ? h1 : l1; slot[1] = h ? 7 : 1;
        out[2] = h ? l0 : h0; slot[2] = h ? 2 : 4;
        out[3] = h ? l1 : h1; slot[3] = h ? 3 : 5;
    }
};

// F16: the row already holds fp16 values in natural K order — weight types without a converter of their own are dequantized into an fp16
// scratch matrix first (abi.cu: b200_launch_dequant_f16) and then take this path.  Thread half h owns K elements [32 h, 32 h + 32); every
// group of four goes out in the activations' (0, 2, 1, 3) order.
template <> struct w_frag<B200_TYPE_F16> {
    uint4 q[4];
    __device__ __forceinline__ void load(const uint32_t row, const int, const int h) {
#pragma unroll
        for (int c = 0; c < 4; ++c) q[c] = lds_u4(row + 64 * h + 16 * c);
    }
    static __device__ __forceinline__ void load_row(w_frag & f0, w_frag & f1, const uint32_t row, const int sub) { f0.load(row, sub, 0); f1.load(row, sub, 1); }
    __device__ __forceinline__ void convert(const int, const int h, const uint32_t, uint4 (&out)[4], int (&slot)[4]) const {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            out[c] = make_uint4(__byte_perm(q[c].x, q[c].y, 0x5410), __byte_perm(q[c].x, q[c].y, 0x7632),
                                __byte_perm(q[c].z, q[c].w, 0x5410), __byte_perm(q[c].z, q[c].w, 0x7632));
            slot[c] = 4 * h + c;
        }
    }
};

// Q4_0 (ggml-common.h:184-189, dequantize_row_q4_0 ggml-quants.c:314-332): stage = blocks 2it, 2it+1; thread half h owns
// block 2it+h: low nibbles = elements 0..15, high nibbles = elements 16..31, value (q - 8) * d.
template <> struct w_frag<B200_TYPE_Q4_0> {
    uint32_t w[5];
    // row = shared-memory address of this row's 8 blocks; sub = pair of blocks inside them
    __device__ __forceinline__ void load(const uint32_t row, const int sub, const int h) {
        const uint32_t b = row + (2 * sub + h) * SZ_Q4_0 - 2 * h;              // 4-byte aligned (18 * odd - 2)
#pragma unroll
        for (int j = 0; j < 5; ++j) w[j] = lds_u32(b + 4 * j);
    }
    static __device__ __forceinline__ void load_row(w_frag & f0, w_frag & f1, const uint32_t row, const int sub) { f0.load(row, sub, 0); f1.load(row, sub, 1); }
    __device__ __forceinline__ void convert(const int, const int h, const uint32_t k6400, uint4 (&out)[4], int (&slot)[4]) const {
        const __half2 d2 = __half2half2(__ushort_as_half((unsigned short) (h ? w[0] >> 16 : w[0] & 0xFFFF)));
        const __half2 k1032 = h2_const(1032.0f), sixteenth = h2_const(0.0625f), m72 = h2_const(-72.0f);
        uint32_t lo[8], hi[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t qs = h ? w[j + 1] : __funnelshift_r(w[j], w[j + 1], 16);
            __half2 a, b, c, d;
            nib_pairs(qs, k6400, a, b, c, d);
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
    // row = shared-memory address of this row's 4 blocks (+ 8 bytes of padding); sub = pair of blocks inside them
    __device__ __forceinline__ void load(const uint32_t row, const int sub, const int h) {
        const uint32_t b = row + (2 * sub + h) * SZ_Q8_0 - 2 * h;              // 4-byte aligned (34 * odd - 2)
#pragma unroll
        for (int j = 0; j < 9; ++j) w[j] = lds_u32(b + 4 * j);
    }
    static __device__ __forceinline__ void load_row(w_frag & f0, w_frag & f1, const uint32_t row, const int sub) { f0.load(row, sub, 0); f1.load(row, sub, 1); }
    __device__ __forceinline__ void convert(const int, const int h, const uint32_t k6400, uint4 (&out)[4], int (&slot)[4]) const {
        (void) k6400;
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
// Q6_K (ggml-common.h:352-358, dequantize_row_q6_K ggml-quants.c:1877-1906): per 128-half n, element l of 32-group r is
// (nibble r>>1 of ql[64n + 32(r&1) + l]) | (bits 2r, 2r+1 of qh[32n + l]) << 4, value (q - 32) * d * sc[8n + 2r + l/16].
// Stage `sub` = K [64 sub, 64 sub + 64): half n = sub >> 1; thread half h owns group r = 2 (sub & 1) + h.  A 210-byte
// super-block is only 2-byte aligned (in global memory and, after the aligned TMA fetch, in its 224-byte shared-memory row).
template <> struct w_frag<B200_TYPE_Q6_K> {
    uint4 l0, l1, q0, q1;
    uint32_t sc, d;
    // 16 bytes from a 2-byte aligned shared-memory address (the super-block sits `skew` bytes into its 224-byte row)
    static __device__ __forceinline__ uint4 lds16_b2(const uint32_t addr) {
        const uint32_t a = addr & ~3u, sh = (addr & 2u) * 8u;
        const uint32_t w0 = lds_u32(a), w1 = lds_u32(a + 4), w2 = lds_u32(a + 8), w3 = lds_u32(a + 12), w4 = lds_u32(a + 16);
        return make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh), __funnelshift_r(w3, w4, sh));
    }
    __device__ __forceinline__ void load(const uint32_t row, const int sub, const int h) {
        const int n = sub >> 1, r = 2 * (sub & 1) + h;
        l0 = lds16_b2(row + 64 * n + 32 * h);  l1 = lds16_b2(row + 64 * n + 32 * h + 16);
        q0 = lds16_b2(row + 128 + 32 * n);     q1 = lds16_b2(row + 128 + 32 * n + 16);
        asm volatile("ld.shared.u16 %0, [%1];" : "=r"(sc) : "r"(row + 192 + 8 * n + 2 * r));
        asm volatile("ld.shared.u16 %0, [%1];" : "=r"(d)  : "r"(row + 208));
    }
    static __device__ __forceinline__ void load_row(w_frag & f0, w_frag & f1, const uint32_t row, const int sub) { f0.load(row, sub, 0); f1.load(row, sub, 1); }
    __device__ __forceinline__ void convert(const int sub, const int h, const uint32_t k6400, uint4 (&out)[4], int (&slot)[4]) const {
        (void) k6400;
        const int r = 2 * (sub & 1) + h, shl = 4 * (sub & 1), shh = 2 * r;
        const float df = __half2float(__ushort_as_half((unsigned short) d));
        const __half2 s_a = h2_const(df * (float) (int8_t) (sc & 0xff)), s_b = h2_const(df * (float) (int8_t) (sc >> 8));
        const __half2 k1056 = h2_const(1056.0f);
        const uint32_t lw[8] = {l0.x, l0.y, l0.z, l0.w, l1.x, l1.y, l1.z, l1.w};
        const uint32_t hw[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
        uint32_t v[16];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t q = ((lw[j] >> shl) & 0x0F0F0F0Fu) | (((hw[j] >> shh) & 0x03030303u) << 4);      // 4 x (0..63)
            const __half2 a = bits_h2(__byte_perm(q, 0x64646464u, 0x4240));     // (1024 + q0, 1024 + q2)
            const __half2 b = bits_h2(__byte_perm(q, 0x64646464u, 0x4341));     // (1024 + q1, 1024 + q3)
            const __half2 s = j < 4 ? s_a : s_b;
            v[2 * j]     = h2_bits(__hmul2(__hsub2(a, k1056), s));
            v[2 * j + 1] = h2_bits(__hmul2(__hsub2(b, k1056), s));
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            out[c]  = make_uint4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
            slot[c] = 4 * h + c;
        }
    }
};


// ---- the GEMM ---------------------------------------------------------------------------------------------------------
// Persistent: CTA c works on tiles c, c + gridDim.x, ...  MUL_MAT: column tile fastest — the CTAs that run together then work on
// the column tiles of the same few row tiles, so a weight tile is read from HBM once and the other readers hit L2 (the fp16
// activations, a few MB, stay L2-resident anyway).  All roles walk the same tile list and the same running stage counter, so ring
// slots and mbarrier phases carry over from one tile to the next.
struct tc_tile {
    int mat;            // weight matrix index (batch / expert) = 3rd TMA coordinate
    int row0, pos0, nvalid;
    int wsel;           // which weight tensor map (several matrices per launch)
    int m;              // height of the matrix = column stride of its output
    float * Y;
};

// (32-bit arithmetic throughout: the host refuses launches with 2^31 or more work items, and 64-bit divisions here sat in front of
//  the first TMA of every CTA — 1.4 us of a 11 us launch in the phase trace)
// tm = weight rows per work tile (128, or 256 for a CTA pair: CTA `rank` of the pair owns rows [128 rank, 128 rank + 128) of it)
__device__ __forceinline__ bool tc_resolve(const tc_args & a, uint32_t T, tc_tile & t, const int tm = TC_M, const int rank = 0) {
    const uint32_t row_tiles = (uint32_t) (a.n_mats > 1 ? a.rt_end[a.n_mats - 1] : (a.m + tm - 1) / tm);
    const int NT = a.nt;
    int rt, ct = 0, batch = 0;
    if (a.mode == MM_PLAIN) {
        const uint32_t col_tiles = (uint32_t) a.col_tiles;
        if (col_tiles > 1) { ct = (int) (T % col_tiles); T /= col_tiles; }
        rt = (int) (T % row_tiles);
        batch = (int) (T / row_tiles);
    } else {
        rt = (int) (T % row_tiles);
        ct = (int) (T / row_tiles);
    }
    t.wsel = 0; t.m = a.m; t.Y = a.Y;
    if (a.n_mats > 1) {
        const int j = (rt >= a.rt_end[0]) + (rt >= a.rt_end[1]);
        rt -= j == 0 ? 0 : a.rt_end[j - 1];
        t.wsel = j; t.m = a.m_j[j]; t.Y = a.Y_j[j];
    }
    t.row0 = rt * tm + TC_M * rank;
    if (a.mode == MM_PLAIN) {
        int i12 = 0, i13 = 0;
        t.mat = 0;
        if (batch > 0 || a.ne12 > 1) {                   // (a plain 2-D matmul has nothing to divide)
            i12 = (int) ((uint32_t) batch % (uint32_t) a.ne12); i13 = (int) ((uint32_t) batch / (uint32_t) a.ne12);
            t.mat = (int) ((uint32_t) i13 / (uint32_t) a.r3) * a.ne02 + (int) ((uint32_t) i12 / (uint32_t) a.r2);
        }
        t.pos0   = batch * a.n + ct * NT;
        t.nvalid = min(NT, a.n - ct * NT);
        return true;
    }
    int e = 0, lo = 0, hi = 0;
    for (; e < a.n_expert; ++e) {
        lo = __ldg(a.expert_bounds + e);
        hi = __ldg(a.expert_bounds + e + 1);
        const int nt = (hi - lo + NT - 1) / NT;
        if (ct < nt) break;
        ct -= nt;
    }
    if (e >= a.n_expert) return false;
    t.mat    = e;
    t.pos0   = lo + ct * NT;
    t.nvalid = min(NT, hi - t.pos0);
    return true;
}

// What a CTA works on next: a whole tile, or (split K) the K range [k0, k1) of one.
struct tc_seg {
    tc_tile t;
    uint32_t T;
    int k0, k1;             // in 64-wide stages
    int order, nparts;      // position of this part in the tile's fixed summation order
};

struct tc_walker {
    uint32_t w;             // next work item: tile = w % n_tiles, K part = w / n_tiles (the parts of a tile run in the same wave)
    int      stride, tm, rank;

    // pair = false: one CTA per work item; true: the two CTAs of a cluster walk the same items (256-row tiles)
    __device__ __forceinline__ tc_walker(const bool pair = false, const int cta_rank = 0) {
        w = pair ? blockIdx.x >> 1 : blockIdx.x; stride = pair ? gridDim.x >> 1 : gridDim.x; tm = pair ? 2 * TC_M : TC_M; rank = cta_rank;
    }

    __device__ __forceinline__ bool next(const tc_args & a, tc_seg & s) {
        const uint32_t n_tiles = (uint32_t) a.n_tiles, n_items = n_tiles * (uint32_t) a.split_k;
        for (; w < n_items; w += stride) {
            const uint32_t part = a.split_k == 1 ? 0u : w / n_tiles;
            s.T = w - part * n_tiles;
            if (tc_resolve(a, s.T, s.t, tm, rank)) {
                const int len = a.k / TC_KS / a.split_k;
                s.k0 = (int) part * len; s.k1 = s.k0 + len;
                s.order = (int) part; s.nparts = a.split_k;
                w += stride;
                return true;
            }
        }
        return false;
    }
};

#ifdef TC_TRACE_ENABLE     // development aid: per-stage clock64 stamps of CTA 0 (build with -DTC_TRACE_ENABLE, run tools/tc_trace.py)
__device__ long long g_tc_trace[12][128];
#define TC_TRACE(role, gg) do { if (blockIdx.x == 0 && (gg) < 128) g_tc_trace[role][gg] = clock64(); } while (0)
__device__ unsigned long long g_tc_phase[8];          // %globaltimer stamps of CTA 0: where a short GEMM spends its fixed cost
#define TC_PHASE(i, once) do { if (blockIdx.x == 0 && (once)) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); g_tc_phase[i] = t_; } } while (0)
#else
#define TC_TRACE(role, gg) do { } while (0)
#define TC_PHASE(i, once) do { } while (0)
#endif

// PAIR: two CTAs of a cluster (the two SMs of a TPC) share one 256-row tile: each converts its own 128 rows of A into its own tensor
// memory and fetches HALF of the activation tile; one M = 256 MMA (cta_group::2), issued by the leader (cluster rank 0), reads both
// halves — the activation bytes an SM pulls through L2 per flop halve, and that traffic is what bounds single-CTA tiles.
// The barriers "operands ready" (A full, B full) and "accumulator drained" that gate the MMAs live in the LEADER's shared memory, the
// peer's warps arrive there remotely; "slot free" and "accumulator ready" are per CTA, signalled in both by multicast commits.
template <int TYPE, bool PAIR>
__global__ void __launch_bounds__(TC_THREADS, 1) mmq_tc_f16_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmW1,
                                                                   const __grid_constant__ CUtensorMap tmW2, const __grid_constant__ CUtensorMap tmB,
                                                                   const tc_args a) {
    using cfg = tc_cfg<raw_row<TYPE>::value>;
    constexpr int RAW_ROW = raw_row<TYPE>::value, RAW_BYTES = cfg::RAW_BYTES;
    constexpr int BS  = cfg::B_SLOTS, AS = TC_A_SLOTS, RS = cfg::RAW_SLOTS;
    constexpr int SPR = raw_k<TYPE>::value / TC_KS;          // MMA stages per raw slot
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t sB   = base;
    const uint32_t sRaw = sB + BS * TC_B_BYTES;
    const uint32_t sBar = sRaw + RS * RAW_BYTES;
    const uint32_t bBFull = sBar, bBEmpty = bBFull + 8 * BS, bAFull = bBEmpty + 8 * BS, bAEmpty = bAFull + 8 * AS;
    const uint32_t bRFull = bAEmpty + 8 * AS, bREmpty = bRFull + 8 * RS, bAccFull = bREmpty + 8 * RS, bAccEmpty = bAccFull + 16;
    const uint32_t sTmem  = bAccEmpty + 16;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t cta_rank = PAIR ? tc::cluster_ctarank() : 0u;
    constexpr int NC = PAIR ? 2 : 1;     // CTAs whose warps arrive on the leader's barriers
    TC_PHASE(0, tid == 0);

    if (warp == TC_W_TMA) {
        if (lane == 0) {
            tc::tma_prefetch_desc(&tmW);
            tc::tma_prefetch_desc(&tmB);
            // Operands travel to the MMA warp in PAIRS of 64-K stages (one barrier round trip and one pair of commits per 128 K: the issue
            // loop's fixed cost per stage is what an MMA-issue warp among 15 busy warps cannot hide, ncu: the converters spent 41 % of their
            // time waiting for A slots).  Barrier s of a ring stands for slots 2 s and 2 s + 1.
            for (int s = 0; s < BS / 2; ++s) { tc::mbar_init(bBFull + 8 * s, 1); tc::mbar_init(bBEmpty + 8 * s, 1); }   // expect_tx / commit
            for (int s = 0; s < AS / 2; ++s) { tc::mbar_init(bAFull + 8 * s, 8 * NC); tc::mbar_init(bAEmpty + 8 * s, 1); }   // 2 stages x 4 converter warps per CTA / commit
            for (int s = 0; s < RS; ++s) { tc::mbar_init(bRFull + 8 * s, 1); tc::mbar_init(bREmpty + 8 * s, 4 * SPR); }   // expect_tx / 4 converter warps per stage of the slot
            for (int i = 0; i < 2; ++i)  { tc::mbar_init(bAccFull + 8 * i, 1); tc::mbar_init(bAccEmpty + 8 * i, 4 * NC); } // commit / 4 epilogue warps per CTA
            tc::mbar_fence_init();
        }
        __syncwarp();
        if (PAIR) tc::tmem_alloc2(sTmem, 512); else tc::tmem_alloc(sTmem, 512);
    }
    tc::tc_fence_before();
    if (PAIR) tc::cluster_sync(); else __syncthreads();
    tc::tc_fence_after();
    // where this CTA's warps report "A stage written" / "accumulator drained" / "activation bytes landed": the leader's barriers
    const uint32_t rAFull = PAIR ? tc::mapa_u32(bAFull, 0) : bAFull, rAccEmpty = PAIR ? tc::mapa_u32(bAccEmpty, 0) : bAccEmpty;
    const uint32_t rBFull = PAIR ? tc::mapa_u32(bBFull, 0) : bBFull;
    uint32_t tmem;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(sTmem));
    TC_PHASE(1, tid == 0);
    // everything above overlapped the activation-conversion kernel when launched with the PDL attribute; its output is needed from here on
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    // Ring positions are kept as (slot, phase) pairs advanced by hand: the single-thread roles below are latency chains,
    // so no divisions and no 64-bit arithmetic inside the stage loops.
    if (warp < TC_CONV_WARPS) {
        // ===== converters: raw blocks (smem) -> fp16 -> A stage in tensor memory =============================================
        // Three groups of 4 warps take the stages in turn, so three stages are being converted at any time (a stage is a long,
        // mostly serial instruction chain for one warp).  A warp owns the 32 rows of its TMEM lane quadrant (warp index mod 4); a
        // thread converts the 64 K elements of ITS row of the stage (two 32-element halves) and writes them as 32 columns.
        const int grp = warp >> 2, q = warp & 3;
        const uint32_t raw_row_a = sRaw + (uint32_t) ((32 * q + lane) * RAW_ROW);
        const uint32_t ta = tmem + ((uint32_t) (32 * q) << 16) + (uint32_t) TC_A_COL0;
        // gs: running stage counter of this CTA over all its tiles (every role counts the same stages) — ring slots and phases follow
        // from it in closed form, and a group only visits its own stages (gs mod 3 == grp).  Segments are whole raw slots (host:
        // k and the split parts are multiples of 256), so the raw-slot counter is gs / SPR.
        uint32_t gs = 0, gmod = 0;
        const uint32_t k6400 = a.k6400;
        int tg = 0; (void) tg;
        tc_walker wk(PAIR, (int) cta_rank);
        tc_seg sg;
        while (wk.next(a, sg)) {
            const int nst = sg.k1 - sg.k0;
            const int first = (int) ((grp + 3u - gmod) % 3u);
            for (int lt = first; lt < nst; lt += TC_CONV_GROUPS) {
                const uint32_t g = gs + (uint32_t) lt;
                const int it = sg.k0 + lt, sub = lt & (SPR - 1), sbi = it / SPR;
                const int as = (int) (g & (AS - 1)), ap = as >> 1;            // TMEM slot of the stage, barrier of its pair
                const uint32_t a_ph = (g / AS) & 1u;
                const uint32_t slotc = g / SPR;
                const int rs = (int) (slotc % RS);
                const uint32_t r_ph = (slotc / RS) & 1u;
                // TMA needs 16-byte aligned global addresses: Q8_0's odd raw slots (136-byte steps) are fetched 8 bytes early, Q6_K's
                // slot i (210-byte steps) 2 (i mod 8) bytes early; the converters skip the skew
                const uint32_t skew = raw_skew<TYPE>(sbi);
                tc::mbar_wait(bRFull + 8 * rs, r_ph, 40);
                TC_PHASE(3, tid == 0 && g == 0);
                if (tid == 0) TC_TRACE(6, (int) g);
                w_frag<TYPE> f0, f1;
                uint4 o0[4], o1[4]; int s0[4], s1[4];
                w_frag<TYPE>::load_row(f0, f1, raw_row_a + rs * RAW_BYTES + skew, sub);
                f0.convert(sub, 0, k6400, o0, s0);
                f1.convert(sub, 1, k6400, o1, s1);
                // chunk s of the row (K elements 8 s .. 8 s + 7 of the stage) = columns 4 s .. 4 s + 3 (the slots are
                // compile-time constants once the half is)
                uint32_t v[32];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    v[4 * s0[c]] = o0[c].x; v[4 * s0[c] + 1] = o0[c].y; v[4 * s0[c] + 2] = o0[c].z; v[4 * s0[c] + 3] = o0[c].w;
                    v[4 * s1[c]] = o1[c].x; v[4 * s1[c] + 1] = o1[c].y; v[4 * s1[c] + 2] = o1[c].z; v[4 * s1[c] + 3] = o1[c].w;
                }
                if (tid == 0) TC_TRACE(7, (int) g);
                tc::mbar_wait(bAEmpty + 8 * ap, a_ph ^ 1, 40);
                tc::tc_fence_after();
                if (tid == 0) TC_TRACE(2, (int) g);
                tc::tmem_st32(ta + (uint32_t) (as * TC_A_COLS), v);
                tc::tmem_st_wait();
                tc::tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if (PAIR) tc::mbar_arrive_cluster(rAFull + 8 * ap); else tc::mbar_arrive(bAFull + 8 * ap);
                    tc::mbar_arrive(bREmpty + 8 * rs);
                    if (warp == 0) TC_TRACE(3, (int) g);
                    if (warp == 7) TC_TRACE(5, (int) g);
                }
            }
            gs += (uint32_t) nst;
            gmod = (gmod + (uint32_t) nst) % 3u;
        }
    } else if (warp == TC_W_TMA) {
        // ===== TMA: activation tiles (whole warp in step, one elected lane issues) =========================================
        int bs = 0;
        uint32_t b_ph = 0;
        int tg = 0; (void) tg;
        tc_walker wk(PAIR, (int) cta_rank);
        tc_seg sg;
        while (wk.next(a, sg)) {
            const tc_tile & t = sg.t;
            // only the boxes that hold valid columns; CTA pair: this CTA's half of the N = 16-rounded columns, in 32-column boxes
            const int n16 = (t.nvalid + 15) & ~15, half = n16 >> 1;
            const int nbox = PAIR ? (half + TC_B_BOX_P - 1) / TC_B_BOX_P : (t.nvalid + TC_B_BOX - 1) / TC_B_BOX;
            const int col0 = t.pos0 + (PAIR ? (int) cta_rank * half : 0);
            for (int it = sg.k0; it < sg.k1; it += 2) {                   // (segments are whole pairs of stages: k is a multiple of 128)
                tc::mbar_wait(bBEmpty + 8 * bs, b_ph ^ 1, 100);
                if (lane == 0) TC_TRACE(4, tg);
                tg += 2;
                if (tc::elect_one()) {
                    if (PAIR) {
                        // both halves of the activation tile report to the leader's barrier (it expects the bytes of both)
                        if (cta_rank == 0) tc::mbar_arrive_expect_tx(bBFull + 8 * bs, (uint32_t) (4 * nbox * TC_B_BOX_P_BYTES));
                        for (int h = 0; h < 2; ++h)
                            for (int j = 0; j < nbox; ++j)
                                tc::tma_load_2d_2cta(sB + (2 * bs + h) * TC_B_BYTES + j * TC_B_BOX_P_BYTES, &tmB, rBFull + 8 * bs, (it + h) * TC_KS, col0 + j * TC_B_BOX_P);
                    } else {
                        tc::mbar_arrive_expect_tx(bBFull + 8 * bs, (uint32_t) (2 * nbox * TC_B_BOX_BYTES));
                        for (int h = 0; h < 2; ++h)
                            for (int j = 0; j < nbox; ++j)
                                tc::tma_load_2d(sB + (2 * bs + h) * TC_B_BYTES + j * TC_B_BOX_BYTES, &tmB, bBFull + 8 * bs, (it + h) * TC_KS, col0 + j * TC_B_BOX);
                    }
                }
                __syncwarp();
                if (++bs == BS / 2) { bs = 0; b_ph ^= 1; }
            }
        }
    } else if (warp == TC_W_RAW) {
        // ===== TMA: raw weight blocks, as far ahead as the raw ring allows ==================================================
        int rs = 0;
        uint32_t r_ph = 0;
        int tg = 0; (void) tg;
        tc_walker wk(PAIR, (int) cta_rank);
        tc_seg sg;
        while (wk.next(a, sg)) {
            const tc_tile & t = sg.t;
            const CUtensorMap * tw = t.wsel == 0 ? &tmW : (t.wsel == 1 ? &tmW1 : &tmW2);
            int sbi = sg.k0 / SPR, x = sbi * a.raw_step;                     // next raw slot in the row (16-bit elements)
            for (int it = sg.k0; it < sg.k1; it += SPR) {
                tc::mbar_wait(bREmpty + 8 * rs, r_ph ^ 1, 100);
                if (tc::elect_one()) {
                    tc::mbar_arrive_expect_tx(bRFull + 8 * rs, RAW_BYTES);
                    TC_PHASE(2, tg == 0);
                    tc::tma_load_3d(sRaw + rs * RAW_BYTES, tw, bRFull + 8 * rs, x - (int) (raw_skew<TYPE>(sbi) >> 1), t.row0, t.mat);
                }
                __syncwarp();
                ++tg;
                x += a.raw_step; ++sbi;
                if (++rs == RS) { rs = 0; r_ph ^= 1; }
            }
        }
    } else if (warp == TC_W_MMA) {
        // ===== MMA issue (whole warp in step, one elected lane issues; CTA pair: the leader only) =========================
        if (!PAIR || cta_rank == 0) {
        constexpr uint32_t idesc0 = tc::umma_idesc(/*f32*/ 1, /*f16*/ 0, /*f16*/ 0, PAIR ? 2 * TC_M : TC_M, 0);
        const uint64_t db0 = tc::umma_desc_sw128(sB);
        const uint32_t ta0 = tmem + (uint32_t) TC_A_COL0;
        int as = 0, bs = 0;
        uint32_t a_ph = 0, b_ph = 0, acc = 0, acc_ph = 0;
        int tg = 0; (void) tg;
        tc_walker wk(PAIR, 0);
        tc_seg sg;
        while (wk.next(a, sg)) {
            tc::mbar_wait(bAccEmpty + 8 * acc, acc_ph ^ 1);
            tc::tc_fence_after();
            const uint32_t d_tmem = tmem + acc * TC_NT_MAX;
            const uint32_t idesc = idesc0 | ((uint32_t) ((sg.t.nvalid + 15) >> 4) << 18);      // N = valid columns rounded up to 16 (N >> 3 at bit 17)
            for (int it = sg.k0; it < sg.k1; it += 2) {                   // a pair of stages per round trip
                {   // both operands are normally there: probe the two barriers back to back, wait only for what is missing
                    const bool okA = tc::mbar_try_wait(bAFull + 8 * as, a_ph), okB = tc::mbar_try_wait(bBFull + 8 * bs, b_ph);
                    if (!okA) tc::mbar_wait(bAFull + 8 * as, a_ph);
                    if (!okB) tc::mbar_wait(bBFull + 8 * bs, b_ph);
                }
                if (lane == 0) TC_TRACE(0, tg);
                TC_PHASE(4, lane == 0 && tg == 0);
                tc::tc_fence_after();
                if (lane == 0) TC_TRACE(8, tg);
                const uint32_t da = ta0 + (uint32_t) (2 * as * TC_A_COLS);
                const uint64_t db = db0 + (uint64_t) (2 * bs * (TC_B_BYTES >> 4));
                if (tc::elect_one()) {
                    TC_TRACE(9, tg);
                    if (PAIR) {
#pragma unroll
                        for (int kk = 0; kk < 2 * TC_KS / 16; ++kk)
                            tc::umma_f16_ts_2cta(d_tmem, da + 8 * kk, db + (uint64_t) ((kk >> 2) * (TC_B_BYTES >> 4) + 2 * (kk & 3)), idesc, (uint32_t) (it > sg.k0 || kk > 0));
                        TC_TRACE(10, tg);
                        tc::tc_commit2_mc(bAEmpty + 8 * as, 3);
                        tc::tc_commit2_mc(bBEmpty + 8 * bs, 3);
                    } else {
#pragma unroll
                        for (int kk = 0; kk < 2 * TC_KS / 16; ++kk)
                            tc::umma_f16_ts(d_tmem, da + 8 * kk, db + (uint64_t) ((kk >> 2) * (TC_B_BYTES >> 4) + 2 * (kk & 3)), idesc, (uint32_t) (it > sg.k0 || kk > 0));
                        TC_TRACE(10, tg);
                        tc::tc_commit(bAEmpty + 8 * as);
                        tc::tc_commit(bBEmpty + 8 * bs);
                    }
                    TC_TRACE(11, tg);
                }
                __syncwarp();
                if (lane == 0) TC_TRACE(1, tg);
                tg += 2;
                if (++as == AS / 2) { as = 0; a_ph ^= 1; }
                if (++bs == BS / 2) { bs = 0; b_ph ^= 1; }
            }
            if (tc::elect_one()) { if (PAIR) tc::tc_commit2_mc(bAccFull + 8 * acc, 3); else tc::tc_commit(bAccFull + 8 * acc); }
            __syncwarp();
            acc ^= 1; if (acc == 0) acc_ph ^= 1;
        }
        }
    } else if (warp < TC_W_RAW) {
        // ===== epilogue: TMEM -> registers -> global (coalesced along rows; MUL_MAT_ID scatters columns) ===================
        const int q = warp & 3;
        uint32_t acc = 0, acc_ph = 0;
        tc_walker wk(PAIR, (int) cta_rank);
        tc_seg sg;
        while (wk.next(a, sg)) {
            const tc_tile & t = sg.t;
            tc::mbar_wait(bAccFull + 8 * acc, acc_ph, 400);
            TC_PHASE(5, warp == TC_W_EPI && lane == 0 && g_tc_phase[5] < g_tc_phase[0]);
            tc::tc_fence_after();
            const int row = t.row0 + 32 * q + lane;
            int32_t * turnstile = sg.nparts > 1 ? a.tile_sync + sg.T * 4 + q : nullptr;
            if (sg.order > 0) {
                // split K: wait for the parts that go before this one (rows of this lane quadrant only)
                if (lane == 0) {
                    const long long t0 = clock64();
                    int seen;
                    do {
                        asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(turnstile) : "memory");
                        if (clock64() - t0 > 8000000000LL) __trap();
                    } while (seen < sg.order);
                }
                __syncwarp();
            }
#pragma unroll 1
            for (int c0 = 0; c0 < TC_NT_MAX; c0 += 32) {
                if (c0 >= t.nvalid) break;                            // warp-uniform
                uint32_t v[32];
                tc::tmem_ld32(tmem + ((uint32_t) (32 * q) << 16) + (uint32_t) (acc * TC_NT_MAX + c0), v);
                // this lane fetches the scale / destination column of position c0 + lane; shuffles hand them round
                const int  pl  = t.pos0 + c0 + lane;
                const bool pok = c0 + lane < t.nvalid;
                const float   my_sc = pok ? __ldg(a.colscale + pl) : 0.0f;
                const int32_t my_yc = pok ? (a.ycol ? __ldg(a.ycol + pl) : pl) : 0;
                tc::tmem_ld_wait();
                if (sg.order == 0 && a.n_peers == 0) {
#pragma unroll
                    for (int c = 0; c < 32; ++c) {
                        const float   sc = __shfl_sync(0xffffffffu, my_sc, c);
                        const int32_t yc = __shfl_sync(0xffffffffu, my_yc, c);
                        if (c0 + c < t.nvalid && row < t.m) t.Y[(int64_t) yc * t.m + row] = __uint_as_float(v[c]) * sc;
                    }
                } else if (sg.order == 0) {
                    // row split: the same tile also goes to every rank's full-height output (its own loop: the plain epilogue above
                    // must not carry the peer branch — measured +10 % on shapes whose last wave exposes the epilogue)
#pragma unroll
                    for (int c = 0; c < 32; ++c) {
                        const float   sc = __shfl_sync(0xffffffffu, my_sc, c);
                        const int32_t yc = __shfl_sync(0xffffffffu, my_yc, c);
                        if (c0 + c < t.nvalid && row < t.m) {
                            const float r = __uint_as_float(v[c]) * sc;
                            t.Y[(int64_t) yc * t.m + row] = r;
#pragma unroll 1
                            for (int p = 0; p < a.n_peers; ++p) a.peer_Y[p][(int64_t) yc * a.peer_m + a.peer_row0 + row] = r;
                        }
                    }
                } else {
                    // add to the earlier parts' sum with fire-and-forget reductions (no load round trip); the turnstile orders
                    // the parts, so the summation order — hence the result — is fixed
#pragma unroll
                    for (int c = 0; c < 32; ++c) {
                        const float   sc = __shfl_sync(0xffffffffu, my_sc, c);
                        const int32_t yc = __shfl_sync(0xffffffffu, my_yc, c);
                        if (c0 + c < t.nvalid && row < t.m) {
                            asm volatile("red.relaxed.gpu.global.add.f32 [%0], %1;" :: "l"(t.Y + (int64_t) yc * t.m + row), "f"(__uint_as_float(v[c]) * sc) : "memory");
                        }
                    }
                }
            }
            TC_PHASE(6, warp == TC_W_EPI && lane == 0 && g_tc_phase[6] < g_tc_phase[0]);
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) { if (PAIR) tc::mbar_arrive_cluster(rAccEmpty + 8 * acc); else tc::mbar_arrive(bAccEmpty + 8 * acc); }
            if (turnstile) {
                __threadfence();
                __syncwarp();
                if (lane == 0) asm volatile("red.release.gpu.global.add.s32 [%0], 1;" :: "l"(turnstile) : "memory");
            }
            acc ^= 1; if (acc == 0) acc_ph ^= 1;
        }
    }
    __syncthreads();
    TC_PHASE(7, tid == 0);
#ifdef TC_TRACE_ENABLE
    if (blockIdx.x == 0 && tid == 0) {
        printf("tc phases (ns from entry): set-up done %llu, first weight TMA issued %llu, first raw slot full %llu, first MMA ready %llu, first accumulator full %llu, "
               "first tile stored %llu, all roles done %llu\n", g_tc_phase[1] - g_tc_phase[0], g_tc_phase[2] - g_tc_phase[0], g_tc_phase[3] - g_tc_phase[0],
               g_tc_phase[4] - g_tc_phase[0], g_tc_phase[5] - g_tc_phase[0], g_tc_phase[6] - g_tc_phase[0], g_tc_phase[7] - g_tc_phase[0]);
        const long long t0 = g_tc_trace[0][32];
        for (int gg = 32; gg < 48; ++gg) {
            printf("g=%3d mma_ready=%6lld fenced=%6lld elected=%6lld mmas_issued=%6lld committed=%6lld mma_done=%6lld | conv_start=%6lld converted=%6lld conv_go=%6lld conv_arr0=%6lld conv_arr7=%6lld tma_go=%6lld\n", gg,
                   g_tc_trace[0][gg] - t0, g_tc_trace[8][gg] - t0, g_tc_trace[9][gg] - t0, g_tc_trace[10][gg] - t0, g_tc_trace[11][gg] - t0, g_tc_trace[1][gg] - t0,
                   g_tc_trace[6][gg] - t0, g_tc_trace[7][gg] - t0, g_tc_trace[2][gg] - t0, g_tc_trace[3][gg] - t0, g_tc_trace[5][gg] - t0, g_tc_trace[4][gg] - t0);
        }
    }
#endif
    if (PAIR) tc::cluster_sync();            // the peer may still arrive on this CTA's barriers
    if (warp == TC_W_TMA) {
        __syncwarp();
        tc::tc_fence_after();
        if (PAIR) tc::tmem_dealloc2(tmem, 512); else tc::tmem_dealloc(tmem, 512);
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

bool tc_type_ok(int type) { return type == B200_TYPE_Q4_K || type == B200_TYPE_Q4_0 || type == B200_TYPE_Q8_0 || type == B200_TYPE_Q6_K || type == B200_TYPE_Q5_K || type == B200_TYPE_F16; }

template <int TYPE, bool PAIR>
int launch_one(b200_ctx * ctx, const CUtensorMap * tmw, const CUtensorMap & tm, const tc_args & a) {
    static std::atomic<bool> configured[16];
    auto kern = mmq_tc_f16_kernel<TYPE, PAIR>;
    using cfg_t = tc_cfg<raw_row<TYPE>::value>;
    if (!configured[ctx->device & 15].load(std::memory_order_acquire)) {
        B200_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, cfg_t::SMEM));
        configured[ctx->device & 15].store(true, std::memory_order_release);
    }
    {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned) a.grid); cfg.blockDim = dim3(TC_THREADS);
        cfg.dynamicSmemBytes = cfg_t::SMEM; cfg.stream = ctx->stream;
        cudaLaunchAttribute attr[2];
        int na = 0;
        if (PAIR) {
            attr[na].id = cudaLaunchAttributeClusterDimension;
            attr[na].val.clusterDim.x = 2; attr[na].val.clusterDim.y = 1; attr[na].val.clusterDim.z = 1;
            ++na;
        }
        if (ctx->pdl && ctx->tc_pdl && a.split_k == 1) {      // (split K: a memset sits between the two kernels, nothing to overlap)
            attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[na].val.programmaticStreamSerializationAllowed = 1;
            ++na;
        }
        cfg.attrs = attr; cfg.numAttrs = na;
        B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, tmw[0], tmw[1], tmw[2], tm, a));
    }
    B200_CUDA_TRY(cudaGetLastError());
    ctx->launches++;
    return B200_OK;
}

template <bool PAIR>
int launch_type(b200_ctx * ctx, int type, const CUtensorMap * tmw, const CUtensorMap & tm, const tc_args & a) {
    switch (type) {
        case B200_TYPE_Q4_K: return launch_one<B200_TYPE_Q4_K, PAIR>(ctx, tmw, tm, a);
        case B200_TYPE_Q4_0: return launch_one<B200_TYPE_Q4_0, PAIR>(ctx, tmw, tm, a);
        case B200_TYPE_Q8_0: return launch_one<B200_TYPE_Q8_0, PAIR>(ctx, tmw, tm, a);
        case B200_TYPE_Q6_K: return launch_one<B200_TYPE_Q6_K, PAIR>(ctx, tmw, tm, a);
        case B200_TYPE_Q5_K: return launch_one<B200_TYPE_Q5_K, PAIR>(ctx, tmw, tm, a);
        case B200_TYPE_F16:  return launch_one<B200_TYPE_F16, PAIR>(ctx, tmw, tm, a);
    }
    b200_set_error("tcgen05 GEMM: unsupported weight type %d", type);
    return B200_E_UNSUPPORTED;
}

}  // namespace

int b200_tc_supported(const b200_ctx * ctx, int type, const mm_problem & P) {
    if (!tc_type_ok(type) || !get_encode()) return 0;
    // (fp16 rows: one stage per raw slot, stages travel in pairs — 128; an attention head's K rows are 128 long)
    const int rawk = (type == B200_TYPE_Q8_0 || type == B200_TYPE_F16) ? 128 : 256;
    if (P.k % rawk != 0 || P.k > (1 << 24) || P.m > (1 << 30)) return 0;
    // TMA over the raw blocks: base and strides 16-byte aligned, the matrices of all batches / experts equally spaced
    const uintptr_t bits = (uintptr_t) P.W | (uintptr_t) P.nb01 | (uintptr_t) P.nb02 | (uintptr_t) P.nb03;
    if (bits & 15) return 0;
    if (P.mode == MM_PLAIN) {
        const int64_t ne02 = P.ne12 / P.r2, ne03 = P.ne13 / P.r3;
        if (ne03 > 1 && P.nb03 != P.nb02 * ne02) return 0;
        // (quantized weights: batches are whole matrices one after the other; fp16 matrices may be strided views — the heads of a KV cache
        //  interleave row by row — which the tensor map's strides express as they are)
        if (type != B200_TYPE_F16 && ne02 * ne03 > 1 && P.nb02 < P.nb01 * P.m) return 0;
    } else if (P.n_expert > 1 && P.nb02 < P.nb01 * P.m) return 0;
    if (((uintptr_t) P.X | (uintptr_t) P.nb11 | (uintptr_t) P.nb12 | (uintptr_t) P.nb13) & 15) return 0;
    if (P.mode == MM_ID_PAIRS) return 0;
    if (tc_cfg<>::SMEM > ctx->smem_optin) return 0;
    return 1;
}

struct tc_plan {
    int     nt;          // tile width
    int     split_k;
    bool    pair;        // CTA pairs on 256-row tiles (MM_PLAIN)
    int64_t n, batches, row_tiles, col_tiles, n_tiles;
    size_t  off_xh, off_sc, off_yc, off_sync, total;
};

struct tc_multi {            // several matrices per launch: P.m is the padded total height, these are the parts
    int n_mats;
    const uint8_t * W[3];
    int64_t m[3];
    float * Y[3];
};

// Estimated clocks of one 64-K stage of a tile with w valid columns: the MMA (2 clocks per column of N, N = w rounded up to 16, for a
// single CTA's 128 rows as for a pair's 256) against the bytes one SM pulls through L2 for it — activation boxes + raw weight blocks
// — at the ~42.6 B/clk an SM gets when all 148 pull (profiles/r02_summary.md), whichever is longer.
static double tc_stage_clocks(int64_t w, int raw_stage_bytes, bool pair) {
    const double n16 = (double) ((w + 15) / 16 * 16);
    const double act = pair ? (double) (((int64_t) n16 / 2 + TC_B_BOX_P - 1) / TC_B_BOX_P) * TC_B_BOX_P_BYTES : (double) ((w + TC_B_BOX - 1) / TC_B_BOX) * TC_B_BOX_BYTES;
    const double l2 = (act + raw_stage_bytes) / 42.6;
    return (2.0 * n16 > l2 ? 2.0 * n16 : l2) + 24.0;
}

// longest chain of stages any CTA (or CTA pair) walks when the tiles of a MUL_MAT are dealt round robin (column tile fastest)
static double tc_makespan(int64_t row_tiles, int64_t n, int64_t batches, int nt, int sk, int64_t stages, int units, int raw_stage_bytes, bool pair) {
    const int64_t col_tiles = (n + nt - 1) / nt, n_tiles = row_tiles * col_tiles * batches, items = n_tiles * sk;
    const int64_t last_w = n - (col_tiles - 1) * nt;
    const double c_full = tc_stage_clocks(nt, raw_stage_bytes, pair), c_last = tc_stage_clocks(last_w, raw_stage_bytes, pair);
    const double per_item = (double) (stages / sk) + 24.0;          // + pipeline fill / epilogue hand-over in stage units (measured order of magnitude)
    const int64_t G = items < units ? items : units;
    if (items > 64 * (int64_t) units) {                             // many waves: the average is what counts
        const double avg = (c_full * (col_tiles - 1) + c_last) / (double) col_tiles;
        return avg * per_item * (double) ((items + G - 1) / G);
    }
    double worst = 0.0;
    for (int64_t c = 0; c < G; ++c) {
        double t = 0.0;
        for (int64_t w = c; w < items; w += G) t += ((w % n_tiles) % col_tiles == col_tiles - 1 ? c_last : c_full) * per_item;
        worst = t > worst ? t : worst;
    }
    return worst;
}

static tc_plan make_plan(const b200_ctx * ctx, int type, const mm_problem & P, int64_t n_pos, size_t ws_used, const tc_multi * M = nullptr) {
    tc_plan p;
    p.batches   = P.mode == MM_PLAIN ? P.ne12 * P.ne13 : 1;
    p.n         = P.mode == MM_PLAIN ? P.ne11 : n_pos;
    auto row_tiles_of = [&](int tm) {
        if (!M) return (P.m + tm - 1) / tm;
        int64_t rt = 0;              // every matrix padded to whole tiles (must match tc_args::rt_end)
        for (int j = 0; j < M->n_mats; ++j) rt += (M->m[j] + tm - 1) / tm;
        return rt;
    };
    const int raw_stage_bytes = TC_M * (type == B200_TYPE_Q8_0 ? 68 : (type == B200_TYPE_Q6_K ? 56 : (type == B200_TYPE_Q5_K ? 44 : (type == B200_TYPE_F16 ? 128 : 36))));
    const int forced = ctx->tc_nt >= 192 ? 192 : (ctx->tc_nt >= 128 ? 128 : (ctx->tc_nt > 0 ? 64 : 0));
    p.split_k = 1;
    p.pair = false;
    if (P.mode == MM_PLAIN) {
        // tile width, CTA pairing and split K by the estimated critical path.  ctx->tc_split_k: 0 = never, 1 = auto (long K only: the
        // later parts of a tile are added through a turnstile, which pays off when the K loop is long), 2 / 4 = forced;
        // ctx->tc_pair: 0 = never, 1 = auto, 2 = always
        const int64_t stages = P.k / TC_KS;
        double best = 0.0;
        p.nt = 0;
        for (const int pair : {0, 1}) {
            if ((pair && ctx->tc_pair == 0) || (!pair && ctx->tc_pair == 2)) continue;
            const int64_t rts = row_tiles_of(pair ? 2 * TC_M : TC_M);
            // auto: pairs where they measure faster — matrices with at least two waves of 256-row tiles (14336 x 4096 x 512: 60.6 vs 62.3 us,
            // 128256 x 4096: 449 vs 477 us; shorter grids lose 3-5 % to the cluster launch and the remote barrier hops); never with the
            // fused row-split stores
            if (pair && ctx->tc_pair == 1 && (P.n_peers > 0 || rts * ((p.n + 191) / 192) * p.batches < (int64_t) ctx->sm_count)) continue;
            if (!pair && ctx->tc_pair == 1 && P.n_peers == 0 && row_tiles_of(2 * TC_M) * ((p.n + 191) / 192) * p.batches >= (int64_t) ctx->sm_count) continue;
            const int units = pair ? ctx->sm_count / 2 : ctx->sm_count;
            for (const int nt : {192, 128, 64}) {
                if (forced && nt != forced) continue;
                if (!forced && nt > 64 && p.n <= nt - 64) continue;      // a narrower width already covers n with one tile
                for (const int sk : {1, 2, 4}) {
                    if (sk > 1) {
                        if (pair || ctx->tc_split_k == 0 || P.n_peers > 0 || P.k % (256 * sk) != 0) continue;
                        if (ctx->tc_split_k == 1 && (sk > 2 || P.k < 8192)) continue;
                        if (ctx->tc_split_k > 1 && sk != ctx->tc_split_k) continue;
                    } else if (!pair && ctx->tc_split_k > 1 && P.k % (256 * ctx->tc_split_k) == 0 && P.n_peers == 0) continue;
                    double t = tc_makespan(rts, p.n, p.batches, nt, sk, stages, units, raw_stage_bytes, pair != 0);
                    if (sk > 1) t *= 1.05;                                // the turnstile and the reductions are not free
                    if (p.nt == 0 || t < best) { best = t; p.nt = nt; p.split_k = sk; p.pair = pair != 0; }
                }
            }
        }
        if (p.nt == 0) { p.nt = forced ? forced : 192; p.pair = ctx->tc_pair == 2; }
    } else {
        // MUL_MAT_ID: the positions per expert are only known on the device; take the width that holds an average expert's share
        // with some margin in one tile (a tile costs what its valid columns cost)
        const int64_t avg = n_pos / (P.n_expert > 0 ? P.n_expert : 1);
        p.nt = forced ? forced : (avg <= 48 ? 64 : (avg <= 100 ? 128 : 192));
    }
    p.row_tiles = row_tiles_of(p.pair ? 2 * TC_M : TC_M);
    p.col_tiles = P.mode == MM_PLAIN ? (p.n + p.nt - 1) / p.nt : (n_pos + p.nt - 1) / p.nt + P.n_expert;
    p.n_tiles   = p.row_tiles * p.col_tiles * p.batches;
    auto up = [](size_t v) { return (v + 255) / 256 * 256; };
    p.off_xh   = (ws_used + 1023) / 1024 * 1024;
    p.off_sc   = p.off_xh + up((size_t) n_pos * P.k * sizeof(__half));
    p.off_yc   = p.off_sc + up((size_t) n_pos * 4);
    p.off_sync = p.off_yc + up((size_t) n_pos * 4);
    // (room for the turnstiles of the narrowest tiling: the scratch is reserved before options may change the plan)
    const int64_t max_tiles = row_tiles_of(TC_M) * (P.mode == MM_PLAIN ? (p.n + 63) / 64 : (n_pos + 63) / 64 + P.n_expert) * p.batches;
    p.total    = p.off_sync + up((size_t) max_tiles * 4 * sizeof(int32_t));
    return p;
}

static int launch_gemm_tc_core(b200_ctx * ctx, int type, const mm_problem & P, int64_t n_pos, size_t ws_used, const tc_multi * M);
size_t b200_gemm_tc_scratch(const b200_ctx * ctx, const mm_problem & P, int64_t n_pos, size_t ws_used);

// n_pos = number of activation positions (MM_PLAIN: ne11 * batches; MM_ID_SORTED: pairs).  Uses the context scratch
// AFTER `ws_used` bytes (the caller may keep ids there).
int b200_launch_gemm_tc(b200_ctx * ctx, int type, const mm_problem & P, int64_t n_pos, size_t ws_used) {
    return launch_gemm_tc_core(ctx, type, P, n_pos, ws_used, nullptr);
}

static int launch_gemm_tc_core(b200_ctx * ctx, int type, const mm_problem & P, int64_t n_pos, size_t ws_used, const tc_multi * M) {
    const tc_plan pl = make_plan(ctx, type, P, n_pos, ws_used, M);
    if ((size_t) ctx->ws_size < pl.total) { b200_set_error("tcgen05 GEMM: scratch not reserved"); return B200_E_ARG; }
    uint8_t * ws = (uint8_t *) ctx->ws;
    __half *  Xh = (__half *) (ws + pl.off_xh);
    float *   sc = (float *) (ws + pl.off_sc);
    int32_t * yc = (int32_t *) (ws + pl.off_yc);
    const int64_t n = pl.n;

    {
        cudaLaunchConfig_t cfg = {};
        const bool warp_per_pos = P.k <= 1024 && n_pos >= 1024;
        cfg.gridDim = dim3((unsigned) (warp_per_pos ? (n_pos + 7) / 8 : n_pos)); cfg.blockDim = dim3(256); cfg.stream = ctx->stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr; cfg.numAttrs = (ctx->pdl && ctx->tc_pdl) ? 1 : 0;
        if (warp_per_pos) B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, tc_act_f16_warp_kernel, P, Xh, sc, yc, n_pos));
        else              B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, tc_act_f16_kernel, P, Xh, sc, yc));
    }
    ctx->launches++;

    // fp16 activations: boxes of 64 positions x 64 K (a tile fetches as many as it has valid columns); positions past the end read as zero
    CUtensorMap tmb;
    const cuuint64_t gdim[2]    = {(cuuint64_t) P.k, (cuuint64_t) n_pos};
    const cuuint64_t gstride[1] = {(cuuint64_t) P.k * sizeof(__half)};
    const cuuint32_t box[2]     = {(cuuint32_t) TC_KS, (cuuint32_t) (pl.pair ? TC_B_BOX_P : TC_B_BOX)};
    const cuuint32_t estr[2]    = {1, 1};
    const CUresult cr = get_encode()(&tmb, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, Xh, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) { b200_set_error("cuTensorMapEncodeTiled failed (%d)", (int) cr); return B200_E_CUDA; }

    // raw weight blocks: [nmat][m][row bytes] seen as 16-bit elements, box = 128 rows x 144 bytes
    const int64_t ne02 = P.mode == MM_PLAIN ? P.ne12 / P.r2 : P.n_expert;
    const int64_t nmat = P.mode == MM_PLAIN ? ne02 * (P.ne13 / P.r3) : P.n_expert;
    CUtensorMap tmw[3];
    for (int j = 0; j < 3; ++j) {
        const int jj = M ? (j < M->n_mats ? j : 0) : 0;
        if (j > 0 && jj == 0) { tmw[j] = tmw[0]; continue; }
        const void *  base = M ? (const void *) M->W[jj] : (const void *) P.W;
        const int64_t rows = M ? M->m[jj] : P.m;
        const cuuint64_t wdim[3]    = {(cuuint64_t) (type == B200_TYPE_F16 ? P.k : P.nb01 / 2), (cuuint64_t) rows, (cuuint64_t) (M ? 1 : nmat)};
        const cuuint64_t wstride[2] = {(cuuint64_t) P.nb01, (cuuint64_t) (!M && nmat > 1 ? P.nb02 : P.nb01 * rows)};
        const cuuint32_t wbox[3]    = {(cuuint32_t) raw_row_of(type) / 2, (cuuint32_t) TC_M, 1};
        const cuuint32_t westr[3]   = {1, 1, 1};
        const CUresult wr = get_encode()(&tmw[j], CU_TENSOR_MAP_DATA_TYPE_UINT16, 3, (void *) base, wdim, wstride, wbox, westr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (wr != CUDA_SUCCESS) { b200_set_error("cuTensorMapEncodeTiled (weights) failed (%d)", (int) wr); return B200_E_CUDA; }
    }

    tc_args a = {};
    a.ne02 = (int) ne02;
    a.raw_step = (type == B200_TYPE_Q8_0 ? 4 * SZ_Q8_0 : (type == B200_TYPE_Q6_K ? SZ_Q6_K : (type == B200_TYPE_Q5_K ? SZ_Q5_K : (type == B200_TYPE_F16 ? 2 * TC_KS : TC_RAW_ROW)))) / 2;
    a.W = P.W; a.nb01 = P.nb01; a.nb02 = P.nb02; a.nb03 = P.nb03;
    a.m = (int) P.m; a.k = (int) P.k; a.n = (int) n;
    a.nt = pl.nt; a.col_tiles = (int) pl.col_tiles; a.k6400 = 0x64006400u;
    a.ne12 = (int) P.ne12; a.r2 = (int) P.r2; a.r3 = (int) P.r3;
    a.Y = P.Y; a.colscale = sc; a.ycol = P.mode == MM_PLAIN ? nullptr : yc;
    a.mode = P.mode; a.n_expert = P.n_expert; a.expert_bounds = P.expert_bounds;
    a.n_mats = M ? M->n_mats : 1;
    const int tm = pl.pair ? 2 * TC_M : TC_M;
    if (M) {
        int rt = 0;
        for (int j = 0; j < 3; ++j) {
            const int jj = j < M->n_mats ? j : M->n_mats - 1;
            if (j < M->n_mats) rt += (int) ((M->m[j] + tm - 1) / tm);
            a.rt_end[j] = rt; a.m_j[j] = (int) M->m[jj]; a.Y_j[j] = M->Y[jj];
        }
    }

    if (pl.n_tiles * 4 >= ((int64_t) 1 << 31)) { b200_set_error("tcgen05 GEMM: too many tiles"); return B200_E_UNSUPPORTED; }
    a.n_tiles = pl.n_tiles;
    // Scheduling: whole tiles dealt round robin.  A long-K matrix with few tiles (down projection) is split along K; the parts run
    // in the same wave and the later ones are added to the first through a per-tile turnstile, so the result does not depend
    // on timing.  (Measured and rejected: stream-K with equal runs per CTA — no gain.)
    a.split_k = pl.split_k;
    a.n_peers = P.n_peers; a.peer_m = P.peer_m; a.peer_row0 = P.peer_row0;
    for (int p = 0; p < P.n_peers; ++p) a.peer_Y[p] = P.peer_Y[p];
    const int64_t items = pl.n_tiles * a.split_k;
    a.grid = (int) (items < ctx->sm_count ? items : ctx->sm_count);
    if (pl.pair) { const int64_t cl = ctx->sm_count / 2; a.grid = 2 * (int) (items < cl ? items : cl); }
    if (a.split_k > 1) {
        a.tile_sync = (int32_t *) (ws + pl.off_sync);
        B200_CUDA_TRY(cudaMemsetAsync(a.tile_sync, 0, (size_t) pl.n_tiles * 4 * sizeof(int32_t), ctx->stream));
    }
    return pl.pair ? launch_type<true>(ctx, type, tmw, tmb, a) : launch_type<false>(ctx, type, tmw, tmb, a);
}

// several weight matrices x the same activations (caller checked: same type / k / row size, MM_PLAIN without batch dims)
int b200_mul_mat_multi_tc(b200_ctx * ctx, int n_mats, const b200_tensor * const * a, const b200_tensor * b, const b200_tensor * const * d) {
    mm_problem P = {};
    P.W = (const uint8_t *) a[0]->data; P.nb01 = a[0]->nb[1]; P.nb02 = a[0]->nb[1] * a[0]->ne[1]; P.nb03 = P.nb02;
    P.k = a[0]->ne[0];
    P.X = (const float *) b->data; P.nb11 = b->nb[1]; P.nb12 = b->nb[2]; P.nb13 = b->nb[3];
    P.ne11 = b->ne[1]; P.ne12 = 1; P.ne13 = 1; P.r2 = 1; P.r3 = 1;
    P.Y = (float *) d[0]->data; P.mode = MM_PLAIN;
    tc_multi M = {};
    M.n_mats = n_mats;
    int64_t padded = 0;
    for (int j = 0; j < n_mats; ++j) {
        M.W[j] = (const uint8_t *) a[j]->data; M.m[j] = a[j]->ne[1]; M.Y[j] = (float *) d[j]->data;
        padded += (a[j]->ne[1] + 2 * TC_M - 1) / (2 * TC_M) * (2 * TC_M);      // whole 256-row tiles: valid for single CTAs and for pairs
    }
    P.m = padded;
    if (!b200_tc_supported(ctx, a[0]->type, P)) return B200_E_UNSUPPORTED;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    const int rc = b200_ctx_reserve(ctx, b200_gemm_tc_scratch(ctx, P, b->ne[1], 0));
    if (rc) return rc;
    return launch_gemm_tc_core(ctx, a[0]->type, P, b->ne[1], 0, &M);
}

size_t b200_gemm_tc_scratch(const b200_ctx * ctx, const mm_problem & P, int64_t n_pos, size_t ws_used) { return make_plan(ctx, B200_TYPE_Q4_K, P, n_pos, ws_used).total; }
