/*
 * oracle/ggml_oracle.h — CPU restatement of the reference's quantized MUL_MAT / MUL_MAT_ID path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under llama.cpp-s390x_b200/ (the product) may include, link or
 * call this.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg
 * use it, and only as the checker / the CPU arm — never as the thing shipped.
 *
 * Parity status: PINNED.  Every function here is checked bit-for-bit (integer / byte outputs) or to
 * the last ulp pattern of the same summation order (float outputs) against the reference itself,
 * compiled from /root/reference into oracle/_ref (see oracle/Makefile), by tests/test_oracle.py, both
 * live (when oracle/_ref is present) and against the committed fixtures in tests/golden/.
 *
 * Type ids are the reference's enum ggml_type values (ggml/include/ggml.h:389-432).
 */
#ifndef GGML_ORACLE_H
#define GGML_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum orc_type {
    ORC_TYPE_F32  = 0,
    ORC_TYPE_Q4_0 = 2,
    ORC_TYPE_Q8_0 = 8,
    ORC_TYPE_Q8_1 = 9,
    ORC_TYPE_Q4_K = 12,
    ORC_TYPE_Q6_K = 14,
    ORC_TYPE_Q8_K = 15,
};

/* block geometry (ggml/src/ggml-common.h:184-366) */
int    orc_blck_size(int type);            /* elements per block: 32 or 256                     */
size_t orc_type_size(int type);            /* bytes per block: 18 / 34 / 36 / 144 / 210 / 292    */
size_t orc_row_size(int type, int64_t k);  /* bytes of one row of k elements                    */

/* IEEE binary16 <-> binary32, round-to-nearest-even (ggml/src/ggml-impl.h:384-431) */
float    orc_fp16_to_fp32(uint16_t h);
uint16_t orc_fp32_to_fp16(float f);

/* activation quantizers */
void orc_quantize_row_q8_1(const float * x, void * y, int64_t k);  /* ggml-quants.c:260-293   */
void orc_quantize_row_q8_0(const float * x, void * y, int64_t k);  /* ggml-quants.c:234-257   */
void orc_quantize_row_q8_K(const float * x, void * y, int64_t k);  /* ggml-quants.c:2692-2729 */

/* weight quantizer for the simple type (used to synthesise test weights without the reference) */
void orc_quantize_row_q4_0(const float * x, void * y, int64_t k);  /* ggml-quants.c:71-106    */

/* exact dequantization, type in {Q4_0,Q8_0,Q4_K,Q6_K,Q8_1} (ggml-quants.c:397,491,1467,1877) */
int  orc_dequantize_row(int type, const void * x, float * y, int64_t k);

/* the activation type ggml-cpu pairs with a weight type (ggml-cpu/ggml-cpu.c:226-322):
 * Q4_0,Q8_0 -> Q8_0 ; Q4_K,Q6_K -> Q8_K */
int  orc_vec_dot_type(int wtype);

/* generic vec_dot: one weight row x one quantized activation row (ggml-cpu/quants.c:174-208,
 * 400-428, 645-718, 800-853).  `act` is in orc_vec_dot_type(wtype) format. */
float orc_vec_dot(int wtype, int64_t k, const void * w_row, const void * act);

/*
 * MUL_MAT as ggml-cpu computes it (ggml-cpu/ggml-cpu.c:1151-1439): activations quantized to
 * vec_dot_type, then dst[i0, i1, i2, i3] = vec_dot(W[:, i0, i2/r2, i3/r3], Xq[:, i1, i2, i3]).
 * All tensors contiguous.  W: [k, m, ne02, ne03] blocks; X: [k, n, ne12, ne13] f32;
 * Y: [m, n, ne12, ne13] f32.  Returns 0, or <0 on bad arguments.
 */
int orc_mul_mat(int wtype, const void * W, const float * X, float * Y,
                int64_t m, int64_t n, int64_t k,
                int64_t ne02, int64_t ne03, int64_t ne12, int64_t ne13, int n_threads);

/*
 * MUL_MAT_ID (ggml-cpu/ggml-cpu.c:1443-1694): as: [k, m, n_expert]; b: [k, b_ne1, n_tokens] f32 with
 * b_ne1 in {1, n_used}; ids: n_tokens rows of n_used int32, row stride ids_stride (elements);
 * c: [m, n_used, n_tokens] f32;  c[:, e, t] = as[:, :, ids[e,t]] . b[:, e % b_ne1, t].
 */
int orc_mul_mat_id(int wtype, const void * as, const float * b, const int32_t * ids, float * c,
                   int64_t m, int64_t k, int64_t n_expert, int64_t n_used, int64_t n_tokens,
                   int64_t b_ne1, int64_t ids_stride, int n_threads);

/* f64 ground truth: exact-dequantized W times the *unquantized* f32 X, accumulated in double.
 * Same shapes as orc_mul_mat with ne02=ne03=ne12=ne13=1.  Bounds CPU and GPU error independently. */
int orc_mul_mat_f64(int wtype, const void * W, const float * X, double * Y,
                    int64_t m, int64_t n, int64_t k, int n_threads);

/* NMSE as tests/test-backend-ops.cpp:278-291 : sum((a-b)^2) / sum(a^2) */
double orc_nmse(const float * a, const float * b, size_t n);

#ifdef __cplusplus
}
#endif
#endif
