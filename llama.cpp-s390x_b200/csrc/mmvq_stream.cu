#include <atomic>
// mmvq_stream.cu — the decode path as ONE persistent kernel per list of matmuls ("job table").
//
// Why: a decoder-layer GEMV at n = 1 streams 2-33 MB, i.e. 0.4-5 us of HBM time; launched one kernel per matmul, every one of them
// pays launch -> resolve -> first DRAM round trip -> activation prologue -> tail (measured ~3.6 us of fixed cost per launch, 0.58 of
// the HBM roofline over a Llama-3-8B step).  Here a single grid of one CTA per SM walks the whole list:
//   * every consumer warp owns a private ring of shared-memory slots that an elected lane keeps filled with bulk async copies
//     (cp.async.bulk global -> shared, mbarrier complete_tx) of its NEXT row-pair chunks — across matrix boundaries, so the
//     weight stream never drains between matmuls and up to ~190 KB per SM stay in flight;
//   * four quantizer warps turn the activation vector of the NEXT group of matmuls into Q8_1 values (bit-exact with
//     quantize_row_q8_1_ref, same code as everywhere in this library) in a second shared-memory buffer while the consumers are
//     still on the current one (full / empty mbarriers);
//   * a matmul whose activations are produced by an earlier matmul of the same list waits on a per-job completion counter in
//     global memory (all CTAs done with that job) — only the quantizer warps wait, the weight rings keep filling.
// The arithmetic per row is the one of gemv_q4k (mmvq.cu): same lane mapping, same operation order, same warp reduction, so the
// results are BIT-IDENTICAL to b200_mul_mat and independent of how rows are distributed over CTAs and warps.
//
// Reference behaviour replaced: the vec_dot loop of ggml_compute_forward_mul_mat (/root/reference/ggml/src/ggml-cpu/ggml-cpu.c:1241-1439)
// with ggml_vec_dot_q4_K_q8_K's block algebra (ggml/src/ggml-cpu/quants.c:645-718), for all MUL_MAT nodes of a graph at once.
#include "stream_jobs.cuh"
#include "quantize.cuh"
#include "tc_ptx.cuh"

#include <mutex>

namespace {

#ifndef SQ_NCW_CFG
#define SQ_NCW_CFG 28
#endif
constexpr int SQ_NCW      = SQ_NCW_CFG;                // consumer warps
constexpr int SQ_NQW      = 4;                         // quantizer warps
constexpr int SQ_THREADS  = (SQ_NCW + SQ_NQW) * 32;
constexpr int SQ_CHUNK_SB = 8;                         // super-blocks of one row per chunk (one per 4 lanes)
constexpr int SQ_CHUNK_B  = SQ_CHUNK_SB * SZ_Q4_K;     // 1152
constexpr int SQ_SLOT_B   = 2 * SQ_CHUNK_B;            // a ring slot: the same chunk of two consecutive rows
constexpr int SQ_MAX_S    = 3;
constexpr int SQ_NCNT     = 256;                       // CTA-level row counters, indexed by job % SQ_NCNT (see SQ_MAX_GROUP)

// shared memory: [barriers | progress words | slot meta | row counters | ring of activation images (2 x the largest image) | weight rings]
constexpr int OFF_FULL  = 0;                                         // [SQ_NCW][SQ_MAX_S] u64
constexpr int SQ_NXF    = 16;                                        // image barriers: group g completes xfull[g % 16], phase (g / 16) & 1
constexpr int OFF_XFULL = OFF_FULL + SQ_NCW * SQ_MAX_S * 8;          // [SQ_NXF] u64
constexpr int OFF_GPROG = OFF_XFULL + SQ_NXF * 8;                    // [32] u32: consumer warp w has finished every group < s_gprog[w]
constexpr int OFF_EPOCH = OFF_GPROG + 128;                           // u32: this launch's epoch (+ pad)
constexpr int OFF_META  = OFF_EPOCH + 16;                            // [SQ_NCW][SQ_MAX_S] uint2
constexpr int OFF_CNT   = OFF_META + SQ_NCW * SQ_MAX_S * 8;          // [SQ_NCNT] u32
constexpr int OFF_X     = (OFF_CNT + SQ_NCNT * 4 + 127) / 128 * 128;

__host__ __device__ constexpr int sq_xbytes(int kmax) { return kmax + kmax / 4; }

__device__ __forceinline__ int act_slot(int t) { return t ^ (((t >> 3) & 1) << 1) ^ ((t >> 4) & 1); }     // as in mmvq.cu

__device__ __forceinline__ int dp4a_us(uint32_t a_unsigned, int b_signed, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a_unsigned), "r"(b_signed), "r"(c));
    return d;
}

__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void * src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 r;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
    return r;
}
__device__ __forceinline__ float4 ld_cg_f4(const float * p) {        // L2-coherent: X may have been written by other SMs during this launch
    float4 r;
    asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

struct act_frag { uint4 a0, a1, a2, a3; float4 e; };

// one lane's share of one super-block of one row: 64 weights (sub-blocks 2g, 2g+1) — the operation sequence of q4k_task (mmvq.cu)
__device__ __forceinline__ float q4k_dot(const uint4 hdr, const uint4 q0, const uint4 q1, const int g, const act_frag & A, float r) {
    const int sh = 16 * (g & 1);
    const uint32_t a = hdr.y >> sh, b = hdr.z >> sh, c = hdr.w >> sh;
    uint32_t scp, mnp;
    if (g < 2) {
        scp = a & 0x3f3f;
        mnp = b & 0x3f3f;
    } else {
        scp = (c & 0x0f0f) | ((a & 0xc0c0) >> 2);
        mnp = ((c >> 4) & 0x0f0f) | ((b & 0xc0c0) >> 2);
    }
    const float2 dm  = __half22float2(*reinterpret_cast<const half2 *>(&hdr.x));
    const float dsc0 = dm.x * (float) (scp & 0xff);
    const float dsc1 = dm.x * (float) (scp >> 8) * 0.0625f;          // high nibbles enter the dot product as 16*q
    const float dmn0 = dm.y * (float) (mnp & 0xff);
    const float dmn1 = dm.y * (float) (mnp >> 8);
    const uint32_t m_lo = 0x0F0F0F0Fu, m_hi = 0xF0F0F0F0u;
    int s0 = __dp4a((int) (q0.x & m_lo), (int) A.a0.x, 0);
    s0 = __dp4a((int) (q0.y & m_lo), (int) A.a0.y, s0); s0 = __dp4a((int) (q0.z & m_lo), (int) A.a0.z, s0); s0 = __dp4a((int) (q0.w & m_lo), (int) A.a0.w, s0);
    s0 = __dp4a((int) (q1.x & m_lo), (int) A.a1.x, s0); s0 = __dp4a((int) (q1.y & m_lo), (int) A.a1.y, s0);
    s0 = __dp4a((int) (q1.z & m_lo), (int) A.a1.z, s0); s0 = __dp4a((int) (q1.w & m_lo), (int) A.a1.w, s0);
    int s1 = dp4a_us(q0.x & m_hi, (int) A.a2.x, 0);
    s1 = dp4a_us(q0.y & m_hi, (int) A.a2.y, s1); s1 = dp4a_us(q0.z & m_hi, (int) A.a2.z, s1); s1 = dp4a_us(q0.w & m_hi, (int) A.a2.w, s1);
    s1 = dp4a_us(q1.x & m_hi, (int) A.a3.x, s1); s1 = dp4a_us(q1.y & m_hi, (int) A.a3.y, s1);
    s1 = dp4a_us(q1.z & m_hi, (int) A.a3.z, s1); s1 = dp4a_us(q1.w & m_hi, (int) A.a3.w, s1);
    r = fmaf(dsc0 * A.e.x, (float) s0, r);
    r = fmaf(dsc1 * A.e.z, (float) s1, r);
    r = fmaf(-dmn0, A.e.y, r);
    r = fmaf(-dmn1, A.e.w, r);
    return r;
}

// Activation quantization is done ONCE per group for the whole GPU, not once per CTA: the vector is cut into units (4 .. 32
// blocks of 32 values, at most 64 units), unit u of group g is quantized by a quantizer warp of CTA (37 g + u) mod grid into an
// IMAGE of the shared-memory layout in global memory (L2-resident: 5-18 KB) and announced by a flag word carrying the launch
// epoch; every CTA's coordinator warp waits for the group's flags and pulls the image into its activation buffer with two bulk
// async copies.  (Quantizing in every CTA — 148 times the same work — cost 20 % of the instruction issue of a Llama-3-8B step on
// one GPU and did not shrink with the row split: 0.53 ms per step at 8 GPUs against 0.08 ms of HBM time.)
__host__ __device__ __forceinline__ int sq_unit_blocks(const int nbk) { return nbk <= 256 ? 4 : nbk <= 512 ? 8 : nbk <= 1024 ? 16 : 32; }

// four consecutive activations: from the tensor, or from the exchange entries of the job that produces them — polled until every
// word carries this launch's epoch (the arrival of the word IS the arrival of the value: one 8-byte access, no flag, no fence)
__device__ __forceinline__ float4 sq_load_x4(const float * x, const unsigned long long * ll, const uint32_t epoch, const int idx) {
    if (!ll) return ld_cg_f4(x + idx);
    const unsigned long long * p = ll + idx;
    unsigned long long w0, w1, w2, w3;
    const long long t0 = clock64();
    while (true) {
        // (system scope: with a row split the words are written by other GPUs through NVLink)
        asm volatile("ld.relaxed.sys.global.v2.b64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(p) : "memory");
        asm volatile("ld.relaxed.sys.global.v2.b64 {%0, %1}, [%2];" : "=l"(w2), "=l"(w3) : "l"(p + 2) : "memory");
        if ((uint32_t) (w0 >> 32) == epoch && (uint32_t) (w1 >> 32) == epoch && (uint32_t) (w2 >> 32) == epoch && (uint32_t) (w3 >> 32) == epoch) break;
        __nanosleep(50);
        if (clock64() - t0 > 20000000000LL) __trap();         // ~10 s: a producer that never delivers must not hang the box for ever
    }
    return make_float4(__uint_as_float((uint32_t) w0), __uint_as_float((uint32_t) w1), __uint_as_float((uint32_t) w2), __uint_as_float((uint32_t) w3));
}

// whole warp: quantize blocks [ub unit, ub (unit + 1)) of x (nbk blocks) into the image (bit-exact Q8_1 values, swizzled like the
// shared-memory buffer: int8 plane of k bytes, then the float2 (d, d * sum) plane)
__device__ __forceinline__ void sq_produce_unit(const float * x, const unsigned long long * ll, const uint32_t epoch, const int nbk, const int ub,
                                                const int unit, uint8_t * img, const int lane) {
    const int sub = lane & 7;
    float2 * img_ds = reinterpret_cast<float2 *>(img + (size_t) nbk * 32);
    for (int r0 = 0; r0 < ub; r0 += 8) {
        float4 v[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int ib = unit * ub + r0 + u * 4 + (lane >> 3);
            // (blocks past the unit or the vector: re-read the unit's first block instead of somebody else's, the result is dropped)
            v[u] = sq_load_x4(x, ll, epoch, (ib < nbk && r0 + u * 4 < ub ? ib : unit * ub) * 32 + sub * 4);
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int  ib = unit * ub + r0 + u * 4 + (lane >> 3);
            const bool ok = ib < nbk && r0 + u * 4 < ub;
            const q8_part qp = quantize_q8_1_part(v[u]);
            if (ok) {
                const int t = ib * 2 + (sub >> 2);
                *reinterpret_cast<uint32_t *>(img + 16 * act_slot(t) + 4 * (sub & 3)) = qp.packed;
                if (sub == 0) img_ds[ib] = make_float2(qp.d, __fmul_rn((float) qp.sum, qp.d));
            }
        }
    }
}

// this launch's epoch: launches that use the exchange buffer are counted in done[SQ_MAX_JOBS + 1] (advanced by the last CTA to
// finish; read after griddepcontrol.wait, i.e. after the previous launch has completed)
__device__ __forceinline__ uint32_t sq_epoch(const uint32_t * done) {
    uint32_t e;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(e) : "l"(done + SQ_MAX_JOBS + 1) : "memory");
    return e + 1;
}

// the sequence of (job, row pair, chunk) steps of one consumer warp.  Rows are walked in PAIRS (2 p, 2 p + 1) — the activations are
// fetched once for two rows and the two results go out as one aligned 16-byte store where they are exchanged — and the PAIRS of a
// matrix are what is dealt to the CTAs: CTA b owns pairs [b base + min(b, rem), ... + base + (b < rem)); the last pair of a matrix
// with an odd number of rows holds one row (flags bit 1).  (Single-row steps for matrices with few rows per CTA were measured
// slower: the kernel is instruction-issue-bound and a single-row step carries the same ring and cursor overhead as a pair step.)
struct sq_cursor {
    int      j;          // current job; == n_jobs: exhausted
    uint32_t pair;       // current pair
    uint32_t phi;        // end of this CTA's pairs of job j
    int      c, nch;     // next chunk, chunks per row
    uint32_t rot;        // pairs of the CTA dealt so far, mod SQ_NCW: pair p of job j goes to warp (rot + p) % SQ_NCW
};

__device__ __forceinline__ uint32_t sq_pairs(const sq_job & J, const uint32_t G) { return J.base * G + J.rem; }      // pairs of the matrix (shard)
__device__ __forceinline__ uint32_t sq_rows(const sq_job & J, const uint32_t G) { return 2 * sq_pairs(J, G) - ((J.flags >> 1) & 1u); }

__device__ __forceinline__ void sq_seek(const sq_params & P, sq_cursor & cu, const uint32_t warp, const uint32_t b) {
    while (cu.j < P.n_jobs) {
        const sq_job & J = P.jobs[cu.j];
        const uint32_t cnt = J.base + (b < J.rem ? 1u : 0u);
        const uint32_t pos = warp >= cu.rot ? warp - cu.rot : warp + SQ_NCW - cu.rot;
        if (pos < cnt) {
            const uint32_t plo = b * J.base + (b < J.rem ? b : J.rem);
            cu.pair = plo + pos;
            cu.phi  = plo + cnt;
            cu.c    = 0;
            cu.nch  = (J.nsb + SQ_CHUNK_SB - 1) / SQ_CHUNK_SB;
            return;
        }
        cu.rot = (cu.rot + cnt) % SQ_NCW;
        cu.j++;
    }
}

__device__ __forceinline__ void sq_advance(const sq_params & P, sq_cursor & cu, const uint32_t warp, const uint32_t b) {
    if (++cu.c < cu.nch) return;
    cu.c = 0;
    cu.pair += SQ_NCW;
    if (cu.pair < cu.phi) return;
    const sq_job & J = P.jobs[cu.j];
    const uint32_t cnt = J.base + (b < J.rem ? 1u : 0u);
    cu.rot = (cu.rot + cnt) % SQ_NCW;
    cu.j++;
    sq_seek(P, cu, warp, b);
}

// MODE 0: one GPU.  MODE 1: row split over `world` processes (one per GPU), see sq_params::ll_peer.  MODE 2 (tests): the same row
// split with ALL ranks inside one launch on one GPU — CTAs [r G, (r + 1) G) are rank r, every rank has its own exchange buffer,
// images, flags, counters and activation / result arena (rank r's copy of a vector of rank 0 lies emu_stride bytes further) and
// its shard is rows [r m, (r + 1) m) of the full matrices the job table points to.  One cooperative launch: its blocks wait on
// one another, which separate launches on one GPU must not do.
template <int MODE>
__global__ void __launch_bounds__(SQ_THREADS, 1)
gemv_q4k_stream(const __grid_constant__ sq_params P) {
    constexpr bool MULTI = MODE != 0;
    const uint32_t G   = MODE == 2 ? gridDim.x / (uint32_t) P.world : gridDim.x;       // CTAs of one rank
    const uint32_t cta = MODE == 2 ? blockIdx.x % G : blockIdx.x;
    const uint32_t er  = MODE == 2 ? blockIdx.x / G : 0u;                               // emulated rank of this CTA
    uint32_t * const done  = P.done  + (MODE == 2 ? (size_t) er * (SQ_MAX_JOBS + 2) : 0);
    uint32_t * const xflag = P.xflag + (MODE == 2 ? (size_t) er * SQ_MAX_JOBS * 64 : 0);
    uint8_t *  const ximg  = P.ximg  + (MODE == 2 ? (size_t) er * P.emu_img_stride : 0);
    const unsigned long long * const ll_local = MODE == 2 ? P.ll_peer[er] : P.ll;
    const size_t arena = MODE == 2 ? (size_t) er * P.emu_stride : 0;                   // bytes between rank 0's and this rank's vectors
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t sbase = tc::smem_u32(smem);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int S = P.stages;

    if (tid == 0) {
        for (int i = 0; i < SQ_NCW * SQ_MAX_S; ++i) tc::mbar_init(sbase + OFF_FULL + 8 * i, 1);
        for (int i = 0; i < SQ_NXF; ++i) tc::mbar_init(sbase + OFF_XFULL + 8 * i, 1);      // the coordinator's expect_tx + the bytes of the image copies
        tc::mbar_fence_init();
    }
    if (tid < SQ_NCNT) reinterpret_cast<uint32_t *>(smem + OFF_CNT)[tid] = 0;
    if (tid < 32) reinterpret_cast<uint32_t *>(smem + OFF_GPROG)[tid] = 0;
    volatile uint32_t * s_gprog = reinterpret_cast<volatile uint32_t *>(smem + OFF_GPROG);
    volatile uint32_t * s_epoch = reinterpret_cast<volatile uint32_t *>(smem + OFF_EPOCH);
    __syncthreads();

    if (warp < SQ_NCW) {
        // ================================================= consumer warp ================================================
        const uint32_t bar0  = sbase + OFF_FULL + (uint32_t) warp * SQ_MAX_S * 8;
        const uint32_t meta0 = sbase + OFF_META + (uint32_t) warp * SQ_MAX_S * 8;
        const uint32_t ring0 = sbase + (uint32_t) OFF_X + (uint32_t) P.xring + (uint32_t) (warp * S) * SQ_SLOT_B;
        const int sbl = lane >> 2, g = lane & 3;
        const int xsw = ((g >> 1) << 1) ^ (sbl & 1);
        // lane-constant byte offsets: weights inside a chunk, activations inside a chunk's 8 super-blocks
        const uint32_t w_off = (uint32_t) sbl * SZ_Q4_K;
        const uint32_t q_off = w_off + 16 + 32 * g;
        const uint32_t a_off = (uint32_t) (sbl * 16 + 4 * g) * 16;
        const uint32_t e_off = (uint32_t) (sbl * 8 + 2 * g) * 8;

        sq_cursor pc;                                     // producer cursor: S steps ahead of the consumer
        pc.j = 0; pc.pair = 0; pc.phi = 0; pc.c = 0; pc.nch = 1; pc.rot = 0;
        sq_seek(P, pc, (uint32_t) warp, cta);
        auto fill = [&](const int slot) {
            const sq_job & J = P.jobs[pc.j];
            const int      left  = (int) J.nsb - pc.c * SQ_CHUNK_SB;
            const uint32_t bytes = (uint32_t) (left < SQ_CHUNK_SB ? left : SQ_CHUNK_SB) * SZ_Q4_K;
            const uint32_t prow  = 2 * pc.pair;
            const bool     two   = prow + 1 < sq_rows(J, G);
            const bool     last  = pc.c + 1 == pc.nch;
            const uint8_t * src  = J.W + ((size_t) prow + (MODE == 2 ? (size_t) er * sq_rows(J, G) : 0)) * J.nb01 + (size_t) pc.c * SQ_CHUNK_B;
            const uint32_t dst   = ring0 + (uint32_t) slot * SQ_SLOT_B;
            const uint32_t bar   = bar0 + 8 * slot;
            if (tc::elect_one()) {
                const uint32_t m1 = (uint32_t) pc.j | ((uint32_t) pc.c << 16) | (two ? 1u << 30 : 0u) | (last ? 1u << 31 : 0u);
                asm volatile("st.shared.v2.u32 [%0], {%1, %2};" :: "r"(meta0 + 8 * slot), "r"(prow), "r"(m1) : "memory");
                tc::mbar_arrive_expect_tx(bar, two ? 2 * bytes : bytes);
                bulk_g2s(dst, src, bytes, bar);
                if (two) bulk_g2s(dst + SQ_CHUNK_B, src + J.nb01, bytes, bar);
            }
            __syncwarp();
            sq_advance(P, pc, (uint32_t) warp, cta);
        };
        int produced = 0, consumed = 0;
        // weights are never written by a kernel of this path: the rings are filled before the dependency wait
        for (int s = 0; s < S && pc.j < P.n_jobs; ++s) { fill(s); produced++; }
        asm volatile("griddepcontrol.wait;" ::: "memory");
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

        int   slot = 0;
        uint32_t phase = 0;
        int   cur_g = -1;
        float accA = 0.0f, accB = 0.0f;
        uint32_t * cnt = reinterpret_cast<uint32_t *>(smem + OFF_CNT);
        // move on to activation group `want`: publish that every earlier group is finished by this warp (its image may be overwritten
        // once all 28 warps say so), wait for the new group's image (it sits at gsm[g] of the image ring)
        uint32_t xb = 0;
        auto enter_group = [&](const int want) {
            while (cur_g < want) {
                cur_g++;
                __syncwarp();                                // every lane has finished reading the old image
                if (lane == 0) s_gprog[warp] = (uint32_t) cur_g;
                if (cur_g < P.n_groups) {
                    tc::mbar_wait(sbase + OFF_XFULL + 8 * (cur_g & (SQ_NXF - 1)), (uint32_t) (cur_g / SQ_NXF) & 1u, 20);
                    xb = sbase + OFF_X + P.gsm[cur_g];
                }
            }
        };
        // Completion accounting (only when a later group of the list waits for an earlier job): a warp "passes" job j once all
        // its rows of the jobs <= j are stored — it walks the list in order — and checks in on a CTA counter; the CTA reports job
        // j as done when its 28 consumer warps and its coordinator (which has then seen the job's activations quantized) have checked in.
        // done[j] == grid therefore means: every row of every job <= j is stored, and nobody will read their activations again.
        // The check-in comes BEFORE the wait for the next group's activations, which may be what is waiting for it.
        int passed = 0;
        auto pass_jobs = [&](const int upto) {
            if (P.n_flagged && lane == 0) {
                for (int jj = passed; jj < upto; ++jj) {
                    if (!(P.jobs[jj].flags & 1)) continue;
                    __threadfence_block();                    // release at CTA scope; the reporting warp's fence below is cumulative
                    if (atomicAdd(cnt + (jj & (SQ_NCNT - 1)), 1u) == (uint32_t) SQ_NCW) {
                        cnt[jj & (SQ_NCNT - 1)] = 0;
                        __threadfence();
                        atomicAdd(done + jj, 1u);
                    }
                }
            }
            passed = upto;
        };
        while (consumed < produced) {
            uint32_t row, m1;                                 // (the slot's description was written when its fill was issued)
            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(row), "=r"(m1) : "r"(meta0 + 8 * slot) : "memory");
            const int  j    = (int) (m1 & 0xffffu);
            const int  c    = (int) ((m1 >> 16) & 0x3fffu);
            const bool two  = (m1 >> 30) & 1u;
            const bool last = (m1 >> 31) & 1u;
            const sq_job & J = P.jobs[j];
            if (j > passed) pass_jobs(j);
            if ((int) J.xg != cur_g) enter_group((int) J.xg);
            tc::mbar_wait(bar0 + 8 * slot, phase);
            const int sb = c * SQ_CHUNK_SB + sbl;
            if (sb < (int) J.nsb) {
                const uint32_t ab = xb + (uint32_t) c * (SQ_CHUNK_SB * 256) + a_off;
                act_frag A;
                A.a0 = lds128(ab + 16 * (0 ^ xsw));
                A.a1 = lds128(ab + 16 * (1 ^ xsw));
                A.a2 = lds128(ab + 16 * (2 ^ xsw));
                A.a3 = lds128(ab + 16 * (3 ^ xsw));
                const uint4 ev = lds128(xb + (uint32_t) J.nsb * 256u + (uint32_t) c * (SQ_CHUNK_SB * 64) + e_off);
                A.e = make_float4(__uint_as_float(ev.x), __uint_as_float(ev.y), __uint_as_float(ev.z), __uint_as_float(ev.w));
                const uint32_t ws = ring0 + (uint32_t) slot * SQ_SLOT_B;
                {
                    const uint4 hdr = lds128(ws + w_off), q0 = lds128(ws + q_off), q1 = lds128(ws + q_off + 16);
                    accA = q4k_dot(hdr, q0, q1, g, A, accA);
                }
                if (two) {
                    const uint4 hdr = lds128(ws + SQ_CHUNK_B + w_off), q0 = lds128(ws + SQ_CHUNK_B + q_off), q1 = lds128(ws + SQ_CHUNK_B + q_off + 16);
                    accB = q4k_dot(hdr, q0, q1, g, A, accB);
                }
            }
            if (last) {
                const float sA = warp_sum(accA);
                const float sB = warp_sum(accB);
                if (MULTI) {
                    // lane p delivers to rank p (the butterfly sum left sA / sB in every lane): one NVLink store per rank and row
                    if (lane < P.world) {
                        const uint32_t epoch = *s_epoch;
                        // (row is even, row0 and ll_off are even: the pair's two words are one aligned 16-byte store — one NVLink
                        // packet instead of two; each word is still its own indivisible {value, epoch})
                        unsigned long long * e = P.ll_peer[lane] + (size_t) (epoch & 1u) * P.ll_cap + J.ll_off + J.row0 + row +
                                                 (MODE == 2 ? (size_t) er * sq_rows(J, G) : 0);
                        const unsigned long long wA = ((unsigned long long) epoch << 32) | __float_as_uint(sA), wB = ((unsigned long long) epoch << 32) | __float_as_uint(sB);
                        if (two) asm volatile("st.relaxed.sys.global.v2.b64 [%0], {%1, %2};" :: "l"(e), "l"(wA), "l"(wB) : "memory");
                        else     asm volatile("st.relaxed.sys.global.b64 [%0], %1;" :: "l"(e), "l"(wA) : "memory");
                    }
                } else if (lane == 0) {
                    J.Y[row] = sA;
                    if (two) J.Y[row + 1] = sB;
                    if (J.ll_off != SQ_NO_LL) {
                        unsigned long long * e = P.ll + J.ll_off + row;
                        const uint32_t epoch = *s_epoch;
                        asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" :: "l"(e), "l"(((unsigned long long) epoch << 32) | __float_as_uint(sA)) : "memory");
                        if (two) asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" :: "l"(e + 1), "l"(((unsigned long long) epoch << 32) | __float_as_uint(sB)) : "memory");
                    }
                }
                accA = 0.0f; accB = 0.0f;
            }
            // the slot has been read (values are in registers): refill it with the step S ahead
            __syncwarp();
            if (pc.j < P.n_jobs) { fill(slot); produced++; }
            consumed++;
            if (++slot == S) { slot = 0; phase ^= 1u; }
        }
        pass_jobs(P.n_jobs);
        enter_group(P.n_groups);                              // pass the remaining groups so that the quantizers can finish
    } else {
        // ================================================= quantizer warps ==============================================
        asm volatile("griddepcontrol.wait;" ::: "memory");
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
        const int qw = warp - SQ_NCW;                         // 0: coordinator of this CTA's activation buffers; 1 .. 3: producers of units
        const uint32_t epoch = sq_epoch(done);
        if (qw == 0) {
            if (lane == 0) *s_epoch = epoch;
            unsigned long long * tr = nullptr;
            if (P.trace && lane == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) tr = P.trace + (blockIdx.x == 0 ? 0 : SQ_MAX_JOBS * 4);
            auto stamp = [&](const int g, const int what) {
                if (tr) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); tr[g * 4 + what] = t; }
            };
            int known = 0;                                    // groups [0, known) have all their units in their images
            const long long t_poll = clock64();
            for (int gidx = 0; gidx < P.n_groups; ++gidx) {
                const int jq = P.gfirst[gidx];
                const sq_job & J = P.jobs[jq];
                const int b = gidx & (SQ_NXF - 1);
                const uint32_t par = (uint32_t) (gidx / SQ_NXF) & 1u;
                const int nbk = (int) J.nsb * 8, ub = sq_unit_blocks(nbk);
                const int nunits = (nbk + ub - 1) / ub;
                stamp(gidx, 0);
                // every unit of the group has been announced with this launch's epoch.  The flags of FOUR groups are fetched per round
                // trip (lane l watches units l and l + 32 of each): one acquire round trip per group was ~1 us x 129 groups, the
                // floor of a step whose matrices are small (row shards)
                while (known <= gidx) {
                    uint32_t seen[4][2];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int gq = known + q < P.n_groups ? known + q : P.n_groups - 1;
                        const int nbq = (int) P.jobs[P.gfirst[gq]].nsb * 8, nuq = (nbq + sq_unit_blocks(nbq) - 1) / sq_unit_blocks(nbq);
                        const uint32_t * f = xflag + (size_t) gq * 64;
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            seen[q][h] = epoch;
                            if (lane + 32 * h < nuq) asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(seen[q][h]) : "l"(f + lane + 32 * h) : "memory");
                        }
                    }
                    uint32_t ok = 0;
#pragma unroll
                    for (int q = 0; q < 4; ++q) ok |= (seen[q][0] == epoch && seen[q][1] == epoch ? 1u : 0u) << q;
                    ok = __reduce_and_sync(0xffffffffu, ok);
                    int adv = 0;
                    while (adv < 4 && ((ok >> adv) & 1u) && known + adv < P.n_groups) ++adv;
                    if (adv) { known += adv; __threadfence(); }                  // (relaxed reads + fence = acquire: the images are visible)
                    else {
                        __nanosleep(40);
                        if (clock64() - t_poll > 20000000000LL) __trap();
                    }
                }
                stamp(gidx, 1);
                // the image's place in the ring is free: every consumer warp has finished the groups that lay there (host-computed:
                // gwait[g] = the latest such group), and the barrier of group g - 16 has been waited for by everybody
                {
                    const int need = (int) P.gwait[gidx] > gidx - SQ_NXF ? (int) P.gwait[gidx] : gidx - SQ_NXF;
                    if (need >= 0) {
                        const long long t0 = clock64();
                        while (true) {
                            const uint32_t mine = lane < SQ_NCW ? s_gprog[lane] : 0xffffffffu;
                            if (__reduce_min_sync(0xffffffffu, mine) > (uint32_t) need) break;
                            __nanosleep(20);
                            if (clock64() - t0 > 20000000000LL) __trap();
                        }
                    }
                }
                if (lane == 0) {
                    asm volatile("fence.proxy.async;" ::: "memory");                           // the image was written with ordinary stores
                    const uint8_t * img = ximg + P.gimg[gidx];
                    const uint32_t xbuf = sbase + OFF_X + P.gsm[gidx], bar = sbase + OFF_XFULL + 8 * b;
                    tc::mbar_arrive_expect_tx(bar, (uint32_t) nbk * 40u);
                    bulk_g2s(xbuf, img, (uint32_t) nbk * 40u, bar);                            // int8 plane + (d, d * sum) plane, contiguous
                }
                stamp(gidx, 2);
                // check-in for the flagged jobs of this group: their activations have been read (quantized) for good
                if (P.n_flagged && lane == 0) {
                    uint32_t * cnt = reinterpret_cast<uint32_t *>(smem + OFF_CNT);
                    for (int jj = jq; jj < P.n_jobs && (int) P.jobs[jj].xg == gidx; ++jj) {
                        if (!(P.jobs[jj].flags & 1)) continue;
                        if (atomicAdd(cnt + (jj & (SQ_NCNT - 1)), 1u) == (uint32_t) SQ_NCW) {
                            cnt[jj & (SQ_NCNT - 1)] = 0;
                            __threadfence();
                            atomicAdd(done + jj, 1u);
                        }
                    }
                }
                if (tr) { tc::mbar_wait(sbase + OFF_XFULL + 8 * b, par, 20); stamp(gidx, 3); }
                __syncwarp();
            }
        } else {
            // producers: this CTA's share of the units, in group order; warp qw takes the groups g with g % 3 == qw - 1
            for (int gidx = qw - 1; gidx < P.n_groups; gidx += SQ_NQW - 1) {
                const sq_job & J = P.jobs[P.gfirst[gidx]];
                const int nbk = (int) J.nsb * 8, ub = sq_unit_blocks(nbk);
                const int nunits = (nbk + ub - 1) / ub;
                const uint32_t first = (cta + G - (uint32_t) ((37 * gidx) % (int) G)) % G;    // (37 g + u) % grid == this CTA
                if ((int) first >= nunits) continue;
                if (J.dep >= 0) {
                    // hazards ordered by completion counters: every CTA has finished job `dep` and everything before it
                    if (lane == 0) {
                        const uint32_t * flag = done + J.dep;
                        uint32_t seen;
                        const long long t0 = clock64();
                        while (true) {
                            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(flag) : "memory");
                            if (seen >= G) break;
                            __nanosleep(40);
                            if (clock64() - t0 > 20000000000LL) __trap();
                        }
                    }
                    __syncwarp();
                }
                // (row split: the slot of this launch's epoch parity; entries are indexed by the row of the FULL matrix)
                const unsigned long long * ll = J.xsrc >= 0 ? ll_local + (MULTI ? (size_t) (epoch & 1u) * P.ll_cap : 0) + P.jobs[J.xsrc].ll_off : nullptr;
                uint8_t * img = ximg + P.gimg[gidx];
                const float * xsrc = reinterpret_cast<const float *>(reinterpret_cast<const char *>(J.X) + arena);
                for (int u = (int) first; u < nunits; u += (int) G) {
                    sq_produce_unit(xsrc, ll, epoch, nbk, ub, u, img, lane);
                    __syncwarp();
                    if (lane == 0) {
                        __threadfence();                          // the unit's image before its flag
                        asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(xflag + (size_t) gidx * 64 + u), "r"(epoch) : "memory");
                    }
                }
            }
        }
    }
    if (MULTI) {
        // Row split: the plain full-height result vectors are written from this rank's exchange buffer once the CTA has finished its
        // own rows — in job order by a fixed thread per entry, so that a vector written by two jobs of the list ends up with the
        // later one's values; nobody in this launch reads the plain vectors of its own jobs (they are polled in the buffer).
        __syncthreads();
        const uint32_t epoch = *s_epoch;
        const unsigned long long * base = ll_local + (size_t) (epoch & 1u) * P.ll_cap;
        for (int j = 0; j < P.n_jobs; ++j) {
            const sq_job & J = P.jobs[j];
            float * yj = reinterpret_cast<float *>(reinterpret_cast<char *>(J.Y) + arena);
            for (uint32_t i = cta * SQ_THREADS + tid; i < J.m_full; i += G * SQ_THREADS) {
                const unsigned long long * p = base + J.ll_off + i;
                unsigned long long w;
                const long long t0 = clock64();
                while (true) {
                    asm volatile("ld.relaxed.sys.global.b64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
                    if ((uint32_t) (w >> 32) == epoch) break;
                    __nanosleep(100);
                    if (clock64() - t0 > 20000000000LL) __trap();     // ~10 s: a rank that never delivers must not hang the others for ever
                }
                yj[i] = __uint_as_float((uint32_t) w);
            }
        }
    }
    {
        // the last CTA to finish leaves the counters at zero for the next launch (nobody polls them any more) and counts the launch
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            if (atomicAdd(done + SQ_MAX_JOBS, 1u) == G - 1) {
                if (P.n_flagged) for (int j = 0; j < P.n_jobs; ++j) done[j] = 0;
                done[SQ_MAX_JOBS + 1] += 1;
                done[SQ_MAX_JOBS] = 0;
            }
        }
    }
}

}  // namespace

// Shared-memory split: `stages` weight-ring slots per consumer warp (2: measured as fast as 3 on the full-size step, 0.788 vs 0.792 ms),
// everything else is the ring of activation images — the deeper it is, the more groups the consumers may be spread over, which is
// what bounds a step of SMALL matrices (row shards): one warp walks the 7 chunks of a k = 14336 row pair in ~5 us, and with two
// image places only ~1.5 decoder layers were in flight (0.27 ms for the Llama-3-8B list at 1/8 of its rows, 3x its HBM time).
int b200_gemv_stream_stages(const b200_ctx * ctx, int64_t kmax) {
    if (kmax <= 0 || kmax % 256 != 0 || kmax > (1 << 20)) return 0;
    const int s = ctx->stream_stages >= 2 && ctx->stream_stages <= SQ_MAX_S ? ctx->stream_stages : 2;
    const size_t rings = (size_t) SQ_NCW * s * SQ_SLOT_B;
    if ((size_t) OFF_X + rings + 2 * (size_t) sq_xbytes((int) kmax) + 256 > ctx->smem_optin) return 0;       // at least two images of the longest vector
    return s;
}

// bytes of the image ring for `stages` weight-ring slots: all the shared memory that is left (a multiple of 128)
size_t b200_gemv_stream_xring(const b200_ctx * ctx, int stages) {
    return (ctx->smem_optin - (size_t) OFF_X - (size_t) SQ_NCW * stages * SQ_SLOT_B) & ~(size_t) 127;
}

int b200_gemv_stream_grid(const b200_ctx * ctx) {
    return ctx->stream_grid > 0 && ctx->stream_grid < ctx->sm_count ? ctx->stream_grid : ctx->sm_count;
}

int b200_launch_gemv_stream(b200_ctx * ctx, sq_params & P) {
    if (P.n_jobs <= 0) return B200_OK;
    if (P.n_jobs > SQ_MAX_JOBS || P.n_groups < 1 || P.n_groups > P.n_jobs) { b200_set_error("gemv_stream: bad job table"); return B200_E_ARG; }
    if (P.stages < 2 || P.stages > SQ_MAX_S || P.xring < 2 * (unsigned) sq_xbytes(P.kmax)) {
        b200_set_error("gemv_stream: k = %d does not fit the shared-memory rings", P.kmax); return B200_E_UNSUPPORTED;
    }
    const size_t smem = (size_t) OFF_X + P.xring + (size_t) SQ_NCW * P.stages * SQ_SLOT_B;
    if (smem > ctx->smem_optin) { b200_set_error("gemv_stream: shared-memory budget exceeded"); return B200_E_UNSUPPORTED; }
    const bool multi = P.world > 1;
    const int  mode  = !multi ? 0 : P.emu_stride ? 2 : 1;
    auto kern = mode == 0 ? gemv_q4k_stream<0> : mode == 1 ? gemv_q4k_stream<1> : gemv_q4k_stream<2>;
    static std::mutex attr_mutex;
    static std::atomic<bool> attr_done[3][64];   // per variant and device (several backends of one device may launch from different threads)
    {
        std::lock_guard<std::mutex> lock(attr_mutex);
        if (!attr_done[mode][ctx->device & 63]) {
            B200_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) ctx->smem_optin));
            attr_done[mode][ctx->device & 63] = true;
        }
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim          = dim3((unsigned) (mode == 2 ? P.world * (ctx->sm_count / P.world) : b200_gemv_stream_grid(ctx)));
    cfg.blockDim         = dim3(SQ_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream           = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    if (mode == 2) { attr[0].id = cudaLaunchAttributeCooperative; attr[0].val.cooperative = 1; }
    cfg.attrs    = attr;
    // (row split: no early start of the next launch — its CTAs would only sit on SMs that another rank's kernel may need when
    // several ranks share one GPU in the tests, and a full-size grid leaves no SM free anyway)
    cfg.numAttrs = mode == 2 ? 1 : ctx->pdl && !multi ? 1 : 0;
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, P));
    ctx->launches++;
    return B200_OK;
}
