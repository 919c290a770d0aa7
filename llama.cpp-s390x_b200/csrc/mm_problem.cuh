// mm_problem.cuh — the argument block every matmul kernel of this path receives, and the in-kernel resolution of
// "which weight matrix, which activation columns, which output columns" for the three addressing modes:
//   MM_PLAIN      : MUL_MAT with batch dims + broadcast            (ggml/src/ggml-cpu/ggml-cpu.c:1171-1207 index math)
//   MM_ID_PAIRS   : MUL_MAT_ID, one (token, slot) pair per problem (expert id read in-kernel; decode-sized n_tokens)
//   MM_ID_SORTED  : MUL_MAT_ID after the expert gather (b200_mm_ids): column tiles never straddle two experts
#pragma once
#include "common.cuh"

enum { MM_PLAIN = 0, MM_ID_PAIRS = 1, MM_ID_SORTED = 2 };

struct mm_problem {
    // weights: [k, m, ne02, ne03] blocks
    const uint8_t * W;
    int64_t nb01, nb02, nb03;
    int64_t m, k;
    // activations, f32 source (fused-quantize kernels) ...
    const float * X;
    int64_t nb11, nb12, nb13;
    // ... and / or pre-quantized planes; plane column of (i1, i2, i3) is i1 + ne11 * (i2 + ne12 * i3)
    q8_planes act;
    int64_t ne11;          // columns per batch (MUL_MAT: n;  MUL_MAT_ID: b.ne[1] in {1, n_used})
    int64_t ne12, ne13;    // batch dims of src1/dst (MUL_MAT_ID: ne12 = n_tokens, ne13 = 1)
    int64_t r2, r3;        // broadcast ratios ne12/ne02, ne13/ne03
    // output: contiguous f32, column stride m
    float * Y;
    // optional addend laid out like Y (a bias or the residual stream): y = W x + bias, added where the result is stored (per-launch GEMV
    // kernels only; single matrix)
    const float * bias;
    // MUL_MAT_ID
    int            mode;
    const int32_t *ids;          // [n_used, n_tokens] with row stride ids_ne_stride (in int32 elements)
    int64_t        ids_stride;
    int            n_used;
    int            n_expert;
    const int32_t *ids_src1;     // MM_ID_SORTED: sorted position -> pair index (token * n_used + slot)
    const int32_t *expert_bounds;// MM_ID_SORTED: [n_expert + 1]
    // row-split all-gather fused into the epilogue: besides Y, every result is also stored into the full-height output
    // of each peer GPU through NVLink peer mappings: peer_Y[p][col * peer_m + peer_row0 + row]   (MM_PLAIN only)
    int     n_peers;
    float * peer_Y[8];
    int64_t peer_m, peer_row0;
    // ... and, with one process per GPU, the completion protocol that replaces a barrier after every matmul: the last CTA
    // of a launch publishes the launch's epoch number into flag[my rank] of every rank; the next launch, before it reads
    // its activations, waits until all ranks' flags have reached the epoch of the launch before it (gemv_q4k).
    uint32_t * sync_flags;        // this rank's flags[world] (peers write them through NVLink)
    uint32_t * peer_flags[8];     // flag arrays of all ranks, indexed like peer_Y
    uint32_t * sync_state;        // local: [0] epoch of the last finished launch, [1] CTAs of the current launch that are done
    int        sync_rank, sync_world;
    // several weight matrices against the SAME activations in one launch (q/k/v, gate/up): rows are numbered through all of
    // them, m = total; matrix j > 0 starts at row m_end[j-1].  Wv / Yv are "virtual bases" (W_j - start_j * nb01, Y_j - start_j),
    // so row r of the numbering is addressed as base + r like in the single-matrix case.  MM_PLAIN, n = 1, no batch dims.
    int             n_mats;
    int64_t         m_end[2];
    const uint8_t * Wv[2];
    float *         Yv[2];
    // row split with the "LL" exchange (one process per GPU, n = 1): instead of plain stores + a completion protocol, every
    // result travels as ONE 8-byte store {value bits, epoch} into slot (ll_off[matrix] + row) of every rank's exchange buffer
    // (ll_dst[p], peer-mapped, own rank included).  An 8-byte store is indivisible, so the arrival of the epoch IS the
    // arrival of the value: no fences, no flags, no barrier — the receiver polls the slot (ll_unpack_kernel).
    // The epoch lives in device memory (ll_state[0] = epoch of the last completed exchange; the receiver advances it), so a
    // CUDA graph that captured these launches uses fresh epochs on every replay.  Slot of epoch e: ll_dst[p] + (e & 1) * ll_cap.
    int              ll_world;
    uint2 *          ll_dst[8];
    int64_t          ll_off[3];
    int64_t          ll_cap;
    const uint32_t * ll_state;
};



// store one result (row, plane-order column index `ycol` = (Y column pointer - P.Y) / m) locally and to the peers
__device__ __forceinline__ void mm_store(const mm_problem & P, float * ycol_ptr, const int64_t row, float v) {
    if (P.bias) v += P.bias[(ycol_ptr - P.Y) + row];
    ycol_ptr[row] = v;
    if (P.n_peers > 0) {
        const int64_t col = (ycol_ptr - P.Y) / P.m;
#pragma unroll 1
        for (int p = 0; p < P.n_peers; ++p) P.peer_Y[p][col * P.peer_m + P.peer_row0 + row] = v;
    }
}

// What one CTA works on: a weight matrix and up to NC columns (plane column index + output pointer each).
template <int NC>
struct mm_cols {
    const uint8_t * W;
    int     nc;                 // live columns (<= NC); dead columns alias column 0 and are never stored
    int64_t plane_col[NC];
    const float * x[NC];        // f32 source column (fused kernels)
    float * y[NC];
};

// tile = blockIdx.y.  Returns false when the tile is empty (MM_ID_SORTED launches an upper bound of tiles).
template <int NC>
__device__ __forceinline__ bool mm_resolve(const mm_problem & P, const int64_t tile, mm_cols<NC> & C) {
    if (P.mode == MM_PLAIN) {
        const int64_t ntiles = (P.ne11 + NC - 1) / NC;
        const int64_t batch  = tile / ntiles;
        const int64_t c0     = (tile - batch * ntiles) * NC;
        const int64_t i12    = batch % P.ne12;
        const int64_t i13    = batch / P.ne12;
        C.W  = P.W + (i12 / P.r2) * P.nb02 + (i13 / P.r3) * P.nb03;
        const int64_t left = P.ne11 - c0;
        C.nc = (int) (left < NC ? left : NC);
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const int64_t i11 = c0 + (c < C.nc ? c : 0);
            C.plane_col[c] = i11 + P.ne11 * batch;
            C.x[c] = reinterpret_cast<const float *>(reinterpret_cast<const char *>(P.X) + i11 * P.nb11 + i12 * P.nb12 + i13 * P.nb13);
            C.y[c] = P.Y + (batch * P.ne11 + i11) * P.m;
        }
        return true;
    }
    if (P.mode == MM_ID_PAIRS) {
        // one pair per tile (NC columns are not used beyond the first)
        const int64_t token = tile / P.n_used;
        const int     slot  = (int) (tile - token * P.n_used);
        const int     e     = __ldg(P.ids + token * P.ids_stride + slot);
        C.W  = P.W + (int64_t) e * P.nb02;
        C.nc = 1;
        const int64_t i11 = slot % P.ne11;
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            C.plane_col[c] = i11 + P.ne11 * token;
            C.x[c] = reinterpret_cast<const float *>(reinterpret_cast<const char *>(P.X) + i11 * P.nb11 + token * P.nb12);
            C.y[c] = P.Y + tile * P.m;
        }
        return e >= 0 && e < P.n_expert;
    }
    // MM_ID_SORTED: walk the experts' tile counts to find (expert, local tile)
    int64_t t = tile;
    int     e = 0;
    int     lo = 0, hi = 0;
    for (; e < P.n_expert; ++e) {
        lo = __ldg(P.expert_bounds + e);
        hi = __ldg(P.expert_bounds + e + 1);
        const int64_t nt = (hi - lo + NC - 1) / NC;
        if (t < nt) break;
        t -= nt;
    }
    if (e >= P.n_expert) return false;
    C.W = P.W + (int64_t) e * P.nb02;
    const int pos0 = lo + (int) t * NC;
    const int left = hi - pos0;
    C.nc = left < NC ? left : NC;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
        const int     pair  = __ldg(P.ids_src1 + pos0 + (c < C.nc ? c : 0));
        const int64_t token = pair / P.n_used;
        const int64_t i11   = (pair - token * P.n_used) % P.ne11;
        C.plane_col[c] = i11 + P.ne11 * token;
        C.x[c] = reinterpret_cast<const float *>(reinterpret_cast<const char *>(P.X) + i11 * P.nb11 + token * P.nb12);
        C.y[c] = P.Y + (int64_t) pair * P.m;
    }
    return true;
}
