/* libsnark_b200.h — C-ABI of the B200 drop-in for libsnark's Groth16 proving hot path.
 *
 * The reference (clearmatics/libsnark + libff + libfqfft) has no FFI on this path: the
 * operators are C++ templates. This header is the boundary a maintainer
 * binds them to; libsnark_b200/overlay/libsnark_b200_overlay.hpp is that binding
 * (template specialisations that forward to these entry points). Conventions follow the
 * in-tree precedent depends/libff/ffi/ffi.h:19-95: extern "C", (const void*, size_t)
 * buffers in the reference's own in-memory layouts, status return, exact-size checks.
 *
 * Layouts (all little-endian u64 limbs, Montgomery form, value < modulus — exactly
 * Fp_model::mont_repr.data, fp.hpp:43 / bigint.hpp:33):
 *   Fr / Fq element      : n limbs                      (alt_bn128 4/4, bls12_377 4/6, bw6_761 6/12)
 *   G1 point             : X | Y | Z                    (3 * nq limbs; alt_bn128_g1.hpp)
 *   G2 point             : X.c0 X.c1 | Y.c0 Y.c1 | Z.c0 Z.c1   (6 * nq limbs; bw6_761: X | Y | Z over Fq)
 *   "special" base       : Z == one (affine) or Z == 0 (infinity)        (alt_bn128_g1.cpp:68-91)
 * MSM results are returned normalised: (x, y, one) or the reference's zero() encoding
 * (0, one, 0), so they memcmp-equal the reference's result after to_affine_coordinates().
 *
 * Every call is synchronous from the caller's view and returns B200_OK (0) or a negative
 * status; b200_last_error() describes the last failure on the calling thread. There is no
 * CPU fallback: without a CUDA device every compute entry point returns B200_ERR_CUDA.
 */
#ifndef LIBSNARK_B200_H_
#define LIBSNARK_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default) /* the library itself is built with -fvisibility=hidden */
#endif

enum b200_curve { B200_ALT_BN128 = 0, B200_BLS12_377 = 1, B200_BW6_761 = 2 };
enum b200_group { B200_G1 = 1, B200_G2 = 2 };
enum b200_status {
    B200_OK = 0,
    B200_ERR_ARG = -1,    /* bad curve / group / size / null pointer */
    B200_ERR_CUDA = -2,   /* CUDA runtime failure (incl. no device) */
    B200_ERR_DOMAIN = -3, /* libfqfft DomainSizeException analogue: size not a power of two <= 2^s */
    B200_ERR_HANDLE = -4  /* unknown handle */
};
/* evaluation_domain operations (libfqfft/evaluation_domain/evaluation_domain.hpp:63-78) */
enum b200_ntt_op { B200_FFT = 0, B200_IFFT = 1, B200_COSET_FFT = 2, B200_ICOSET_FFT = 3 };

/* ---- runtime ------------------------------------------------------------------------ */
int b200_version(void);
/* Binds the calling process to CUDA device `device` (one process per GPU) and creates the
 * library's streams. Idempotent for the same device. */
int b200_init(int device);
int b200_device_count(void);
const char *b200_last_error(void);
/* sizes in bytes of the reference's types for `curve` */
size_t b200_sizeof_fr(int curve);
size_t b200_sizeof_group(int curve, int group);

/* ---- device memory (for callers that keep data resident between calls) -------------- */
int b200_dev_alloc(size_t bytes, void **dev_ptr);
int b200_dev_free(void *dev_ptr);
int b200_memcpy_h2d(void *dev_dst, const void *host_src, size_t bytes);
int b200_memcpy_d2h(void *host_dst, const void *dev_src, size_t bytes);
int b200_sync(void);

/* ---- MSM: libff::multi_exp<GroupT, FieldT, multi_exp_method_BDLO12_signed,
 *           multi_exp_base_form_special>  (multiexp.hpp:63-73, multiexp.tcc:643-688) ------
 * bases   : n points in the reference layout, `base_stride` bytes apart (sizeof(GroupT) for
 *           a std::vector<GroupT>; sizeof(knowledge_commitment<G2,G1>) for a B_query).
 * scalars : n Fr elements, Montgomery form, contiguous.
 * out     : one point, reference layout, normalised.
 * The reference's `chunks` argument has no meaning here (one GPU pass) and is dropped. */
int b200_msm(int curve, int group, const void *bases, size_t base_stride, size_t n,
             const void *scalars, void *out);

/* Device-resident base sets (proving-key queries are uploaded once; SURVEY.md §8b "Ownership"). */
typedef uint64_t b200_bases_t;
int b200_bases_upload(int curve, int group, const void *bases, size_t base_stride, size_t n,
                      b200_bases_t *handle);
/* Optional, once per resident set: precomputes the rows [2^(c w)] P_i, w < W (affine), so that an MSM
 * over the set scatters the digits of ALL windows into one set of 2^(c-1) buckets and pays the bucket
 * reduction once instead of W times; c is then chosen larger (fewer windows, fewer additions). Costs W
 * times the set's device memory (2^20 alt_bn128 G1 points: 64 MiB -> 1 GiB) and one field inversion per
 * table entry, once. The reference's counterpart of trading memory for window work is the fixed-base
 * window table of batch_exp (multiexp.tcc:809-946). Results are unchanged (same group element).
 * b200_bases_table_rows: W after a successful precompute, else 0; b200_bases_table_window_bits: the table's c. */
int b200_bases_precompute(b200_bases_t handle);
int b200_bases_table_rows(b200_bases_t handle);
int b200_bases_table_window_bits(b200_bases_t handle);
int b200_bases_free(b200_bases_t handle);
size_t b200_bases_size(b200_bases_t handle);

/* MSM over bases[first, first+n) of a resident set.
 * scalars_on_device != 0: `scalars` is a device pointer (e.g. the H coefficients left on
 * the device by b200_witness_map_h_dev). */
int b200_msm_resident(b200_bases_t handle, size_t first, size_t n, const void *scalars,
                      int scalars_on_device, void *out);

/* Per-window partial sums instead of the folded point: for multi-GPU sharding by base range
 * (SURVEY.md §8e). out_windows receives b200_msm_num_windows() points of
 * b200_msm_window_bytes() each in the library's extended-Jacobian form
 * (X | Y | ZZ | ZZZ, Montgomery limbs; x = X/ZZ, y = Y/ZZZ) — an opaque blob to gather
 * between ranks. b200_msm_fold_windows folds `num_shards` such arrays (concatenated,
 * all from shards of the same size class n_per_shard) into the final normalised point. */
int b200_msm_num_windows(int curve, size_t n);
int b200_msm_window_bits(int curve, size_t n);
size_t b200_msm_window_bytes(int curve, int group);
int b200_msm_resident_windows(b200_bases_t handle, size_t first, size_t n, const void *scalars,
                              int scalars_on_device, void *out_windows);
/* Same, leaving the window sums in device memory (e.g. a torch CUDA tensor that NCCL gathers). */
int b200_msm_resident_windows_dev(b200_bases_t handle, size_t first, size_t n, const void *scalars,
                                  int scalars_on_device, void *dev_out_windows);
int b200_msm_fold_windows(int curve, int group, size_t n_per_shard, const void *windows,
                          size_t num_shards, void *out);

/* libff::multi_exp_filter_one_zero (multiexp.hpp:83-88, multiexp.tcc:690-757; upstream name
 * multi_exp_with_mixed_addition). Same value as b200_msm: scalars 0 and 1 need no host
 * pre-pass on the device path; the counts the reference prints are returned for log parity
 * (stats[0] = zeros skipped, stats[1] = ones added, stats[2] = the rest; may be NULL). */
int b200_msm_filter_one_zero(int curve, int group, const void *bases, size_t base_stride,
                             size_t n, const void *scalars, void *out, size_t *stats);

/* libsnark::kc_multi_exp_with_mixed_addition<G2, G1, Fr, BDLO12_signed, special>
 * (kc_multiexp.hpp:32-44, kc_multiexp.tcc:31-113) over a sparse_vector of
 * knowledge_commitment<G2,G1> (sparse_vector.hpp:33-74): indices[nnz] (size_t, ascending),
 * values[nnz] (G2 | G1 per entry). Entries with min_idx <= index < max_idx use
 * scalars[index - min_idx]. Returns the pair (out_g2, out_g1), normalised. */
int b200_kc_msm(int curve, const size_t *indices, const void *values, size_t nnz,
                size_t min_idx, size_t max_idx, const void *scalars, size_t num_scalars,
                void *out_g2, void *out_g1, size_t *stats);

/* ---- NTT: libfqfft::basic_radix2_domain<Fr>::{FFT,iFFT,cosetFFT,icosetFFT}
 *      (basic_radix2_domain.tcc:54-88), in place on m = 2^k Fr elements, natural order in
 *      and out. `g` = coset generator (Fr, Montgomery) for the coset ops, else ignored/NULL.
 *      m not a power of two, m < 2 or log2(m) > Fr::s  ->  B200_ERR_DOMAIN. ----------------- */
int b200_ntt(int curve, int op, void *data, size_t m, const void *g);
int b200_ntt_dev(int curve, int op, void *dev_data, size_t m, const void *g);
/* b200_ntt_dev without any device allocation. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it). Same arguments, result and status codes as b200_ntt_dev for basic_radix2_domain sizes
 * m = 2^k, k <= B200_NTT_NOALLOC_MAX_LOG; a larger power of two returns B200_ERR_ARG ("exceeds the static scratch"),
 * other sizes B200_ERR_DOMAIN (the step domains are not served here). The ping-pong buffer, the twiddle table, the
 * coset table and the domain constants live in one file-scope device array of B200_NTT_NOALLOC_SCRATCH_BYTES that
 * every GPU which loads the library holds until the process exits; the call performs no cudaMalloc / cudaFree and no
 * allocating library call (call b200_init once at set-up: it creates the library's streams). All calls share that
 * array and it keeps the last call's tables (rebuilt only when curve, size, generator or coset direction change):
 * calls that overlap on the GPU — two host threads in different execution slots, two processes' worth of streams on
 * one context — race; serialise them. */
#define B200_NTT_NOALLOC_MAX_LOG 16
#define B200_NTT_NOALLOC_SCRATCH_BYTES ((((size_t)5 << (B200_NTT_NOALLOC_MAX_LOG - 1)) + 256) * 48)
int b200_ntt_dev_noalloc(int curve, int op, void *dev_data, size_t m, const void *g);
/* basic_radix2_domain::divide_by_Z_on_coset (basic_radix2_domain.tcc:117-126) with
 * coset generator Fr::multiplicative_generator. */
int b200_divide_by_Z_on_coset(int curve, void *data, size_t m);
/* The domain kinds behind the calls above and below, chosen from m the way libfqfft::get_evaluation_domain does
 * (get_evaluation_domain.tcc:32-70), s = the field's 2-adicity FieldT::s:
 *   m = 2^k, k <= s          basic_radix2_domain     (basic_radix2_domain.tcc)
 *   m = 2^(s+1)              extended_radix2_domain  (extended_radix2_domain.tcc:21-236: the order-2^s subgroup and its
 *                            coset by libff::coset_shift(); two size-2^s transforms plus element-wise steps)
 *   m = 2^k + 2^r, k + 1 <= s   step_radix2_domain   (step_radix2_domain.tcc:20-276)
 * b200_ntt*, b200_divide_by_Z_on_coset, the witness-map calls (whole and chain by chain) and the provers serve all three.
 * b200_set_two_adicity mirrors libff's mutable static FieldT::s / FieldT::root_of_unity (e.g. alt_bn128_init.cpp:53-58):
 * a host that lowers them gets the extended domain at 2^(s'+1); s = 0 restores the field's own value (28 / 47 / 46),
 * with which only alt_bn128 reaches the extended domain at a size a GPU holds (2^29 elements). Roots of unity are always
 * derived from the field's true root, as get_root_of_unity does from a consistently lowered pair. */
int b200_set_two_adicity(int curve, int s);
/* The kind that serves a domain of exactly m elements (valid_for_size of the three classes in the factory's order),
 * B200_DOMAIN_NONE when none does. */
#define B200_DOMAIN_NONE 0
#define B200_DOMAIN_BASIC 1
#define B200_DOMAIN_EXTENDED 2
#define B200_DOMAIN_STEP 3
int b200_domain_kind(int curve, size_t m);

/* The NTT chain of libsnark::r1cs_to_qap_witness_map (r1cs_to_qap.tcc:252-330) for the
 * Groth16 case d1 = d2 = d3 = 0: in = the evaluations aA, aB, aC (m each, host);
 * out_H = coefficients_for_H, m + 1 elements.  All seven transforms and the pointwise
 * (A*B - C) / Z run on the device without intermediate copies. */
int b200_witness_map_h(int curve, const void *aA, const void *aB, const void *aC, size_t m,
                       void *out_H);
/* Same with device buffers; dA, dB, dC are clobbered, dH (m + 1 elements) may alias none. */
int b200_witness_map_h_dev(int curve, void *dA, void *dB, void *dC, size_t m, void *dH);

/* ---- R1CS on the device (r1cs_to_qap.tcc:236-248, :293-296) -------- */
/* The constraint system as ONE CSR over Fr: rows [0, nc) are the a-combinations of the constraints,
 * [nc, 2nc) the b-, [2nc, 3nc) the c-combinations (libsnark::r1cs_constraint, r1cs.hpp:53-70).
 * row_ptr has 3*nc + 1 entries; col[e] = linear_term::index (0 is the constant ONE, k the k-th
 * variable, variable.hpp:96-117); val[e] = linear_term::coeff, Montgomery Fr. Uploaded once per
 * proving key. */
typedef uint64_t b200_r1cs_t;
int b200_r1cs_upload(int curve, size_t num_constraints, size_t num_inputs, size_t num_variables,
                     const uint32_t *row_ptr, const uint32_t *col, const void *val,
                     b200_r1cs_t *handle);
int b200_r1cs_free(b200_r1cs_t handle);
/* aA, aB, aC (m elements each, host) as r1cs_to_qap_witness_map builds them before its first
 * transform (r1cs_to_qap.tcc:236-248): w = (1 | primary | auxiliary), num_variables + 1 elements,
 * host. Includes the input-consistency rows aA[nc + i] = w[i], i <= num_inputs. */
int b200_r1cs_evaluate(b200_r1cs_t handle, const void *w, size_t m, void *aA, void *aB, void *aC);
/* The whole of r1cs_to_qap_witness_map for d1 = d2 = d3 = 0 on the device: d_w (device,
 * num_variables + 1 elements) -> dH = coefficients_for_H (device, m + 1 elements). The MSMs can
 * then consume d_w and dH in place (b200_msm_resident with scalars_on_device = 1). */
int b200_witness_map_r1cs_dev(b200_r1cs_t handle, const void *d_w, size_t m, void *dH);
/* libsnark::r1cs_to_qap_witness_map(cs, primary, auxiliary, d1, d2, d3) with its ZK patch (r1cs_to_qap.tcc:209-344; the
 * sibling provers r1cs_ppzksnark / r1cs_se_ppzksnark call it with random d's, Groth16 with zeros): w = (1 | primary |
 * auxiliary) on the host, d123 = d1, d2, d3 (three Fr, Montgomery, host), out_H = coefficients_for_H =
 * H + d2 coef(A) + d1 coef(B) - d3 + d1 d2 Z (m + 1 elements, host; Z through the domain's add_poly_Z:
 * basic_radix2_domain.tcc:109-115, extended_radix2_domain.tcc:203-214, step_radix2_domain.tcc:235-245). Sparse
 * evaluation, the transforms (eight with a non-zero d1 or d2) and the patch run on the device. */
int b200_witness_map_r1cs_zk(b200_r1cs_t handle, const void *w, size_t m, const void *d123, void *out_H);
/* The two halves of b200_witness_map_r1cs_dev, for provers that run the A, B and C chains on different
 * GPUs (r1cs_to_qap.tcc:252-290 and :293-330), on any of the three radix-2 domain kinds.
 * chain : which = 0/1/2 -> the evaluations of the a-/b-/c-combinations from d_w, iFFT, cosetFFT(g) into
 *         d_out (device, m elements).
 * finish: (dA * dB - dC) / Z on the coset, icosetFFT -> dH (device, m + 1 elements); dA is clobbered. */
int b200_witness_chain_r1cs_dev(b200_r1cs_t handle, const void *d_w, size_t m, int which, void *d_out);
int b200_witness_finish_dev(int curve, void *dA, const void *dB, const void *dC, size_t m, void *dH);
/* d_dst[i] = d_src[d_idx[i]], Fr elements, all pointers device (the scalars of a sparse
 * knowledge-commitment query, kc_multiexp.tcc:52-89). */
int b200_gather_dev(int curve, const void *d_src, const uint32_t *d_idx, size_t n, void *d_dst);

/* ---- the prover in one call (r1cs_gg_ppzksnark_prover, r1cs_gg_ppzksnark.tcc:460-603) ---------- */
/* A proving key assembled from handles that are already on the device: the four queries
 * (r1cs_gg_ppzksnark_proving_key, r1cs_gg_ppzksnark.hpp:74-170; B_query split into its G2 and G1
 * halves, B_indices = sparse_vector::indices or NULL when every index 0..num_variables is
 * present), the constraint system, and the five single points in the reference's layout.
 * The handle does not own the queries or the R1CS. domain_size = the m of
 * get_evaluation_domain(num_constraints + num_inputs + 1), a power of two. */
typedef uint64_t b200_groth16_pk_t;
int b200_groth16_pk_create(int curve, const void *alpha_g1, const void *beta_g1, const void *beta_g2,
                           const void *delta_g1, const void *delta_g2, b200_bases_t A_query,
                           b200_bases_t B_query_g2, b200_bases_t B_query_g1, const uint32_t *B_indices,
                           b200_bases_t H_query, b200_bases_t L_query, b200_r1cs_t r1cs,
                           size_t num_constraints, size_t num_inputs, size_t num_variables,
                           size_t domain_size, b200_groth16_pk_t *handle);
int b200_groth16_pk_free(b200_groth16_pk_t handle);
/* One proof. primary_input (num_inputs) and auxiliary_input (num_variables - num_inputs) are host
 * arrays of Montgomery Fr, r and s the zero-knowledge randomness (the reference draws them at
 * .tcc:499-500; the caller supplies them here so that proofs are reproducible). proof_out =
 * g_A (G1) | g_B (G2) | g_C (G1), normalised, reference layouts. stage_ms (nullable, 7 doubles):
 * witness upload, witness map, A, B, H, L, total — the five MSMs run concurrently, so A..L are the wall clock of
 * each inside its own execution slot and do not add up to the total. An auxiliary_input in page-locked memory
 * (cudaHostAlloc / cudaHostRegister) is uploaded from where it lies; pageable memory is staged by a few host threads. */
int b200_groth16_prove(b200_groth16_pk_t handle, const void *primary_input,
                       const void *auxiliary_input, const void *r, const void *s, void *proof_out,
                       double *stage_ms);
/* The same prover with the witness already in HBM: b200_groth16_witness_upload places (1 | primary | auxiliary)
 * in the key's device buffer (what b200_groth16_prove does first), b200_groth16_prove_resident runs everything
 * after that copy — r1cs_gg_ppzksnark.tcc:489-598 on the resident witness — so that a caller who proves several
 * (r, s) over one witness, or a benchmark that separates the PCIe copy from the compute, pays the upload once.
 * stage_ms as b200_groth16_prove with the upload entry 0. */
int b200_groth16_witness_upload(b200_groth16_pk_t handle, const void *primary_input, const void *auxiliary_input);
int b200_groth16_prove_resident(b200_groth16_pk_t handle, const void *r, const void *s, void *proof_out,
                                double *stage_ms);

/* ---- the prover sharded over the GPUs of one box (the reference splits every multi_exp
 *      over OpenMP threads the same way, multiexp.tcc:663-687). One process per GPU; rank g of `world`
 *      holds entries b200_shard_range(total, world, g) of each query: A (num_variables + 1 points),
 *      B (B_total entries of the sparse vector), H (domain_size - 1), L (num_variables - num_inputs),
 *      and the whole constraint system. A rank computes the window sums of its five partial MSMs into
 *      one "blob" of b200_groth16_shard_bytes() bytes in device memory; the ranks all-gather the blobs
 *      (NCCL over NVLink) and b200_groth16_assemble — host code, no device needed — folds them into the
 *      proof. ------------------------------------------------------------------------------------- */
void b200_shard_range(size_t total, int world, int rank, size_t *first, size_t *count);
/* query_sizes = {A, B, H, L} totals as above. */
size_t b200_groth16_shard_bytes(int curve, const size_t query_sizes[4], int world);
/* As b200_groth16_pk_create with the handles holding this rank's shards. B_indices_shard = this rank's slice
 * of sparse_vector::indices (NULL when B_total == num_variables + 1, i.e. every index is present). */
int b200_groth16_pk_create_shard(int curve, const void *alpha_g1, const void *beta_g1, const void *beta_g2,
                                 const void *delta_g1, const void *delta_g2, b200_bases_t A_shard,
                                 b200_bases_t B_g2_shard, b200_bases_t B_g1_shard,
                                 const uint32_t *B_indices_shard, size_t B_total, b200_bases_t H_shard,
                                 b200_bases_t L_shard, b200_r1cs_t r1cs, size_t num_constraints,
                                 size_t num_inputs, size_t num_variables, size_t domain_size, int rank,
                                 int world, b200_groth16_pk_t *handle);
/* One rank's share of one proof in one call: witness (host) in, blob (device) out; the witness map is
 * computed on every rank. stage_ms as b200_groth16_prove. */
int b200_groth16_prove_shard(b200_groth16_pk_t handle, const void *primary_input,
                             const void *auxiliary_input, void *dev_blob, double *stage_ms);
/* The same in steps, so that the caller can place collectives between them:
 *   begin : (1 | primary | auxiliary) to the device.
 *   chain : b200_witness_chain_r1cs_dev on the uploaded witness (which = 0/1/2 for A/B/C) -> dev_out.
 *   h     : coefficients_for_H — from the witness alone when dA = dB = dC = NULL (all seven transforms on
 *           this rank), else b200_witness_finish_dev on three chain results (broadcast between ranks).
 *   msm   : the partial MSMs selected by mask (bit 0 A, 1 B.g2, 2 B.g1, 3 H, 4 L) into their blob slots;
 *           H needs `h` first, the others only `begin`. */
int b200_groth16_shard_begin(b200_groth16_pk_t handle, const void *primary_input, const void *auxiliary_input);
/* The uploaded witness (1 | primary | auxiliary), (num_variables + 1) Fr, copied device-to-device between the key's
 * buffer and dev_buf: to_key = 0 after `begin` on the rank that holds the witness, to_key = 1 on the ranks that
 * received dev_buf by an NCCL broadcast — one host upload per proof instead of one per rank. */
int b200_groth16_shard_witness_copy(b200_groth16_pk_t handle, void *dev_buf, int to_key);
int b200_groth16_shard_chain(b200_groth16_pk_t handle, int which, void *dev_out);
/* coefficients_for_H [first, first + count) copied device-to-device between the key's buffer and dev_buf (count Fr elements):
 * to_key = 0 on the rank that ran `h`, to hand each rank the slice its H shard MSM reads (b200_shard_range of the H query)
 * instead of every rank repeating the pointwise step and the inverse transform; to_key = 1 on the receiving ranks. */
int b200_groth16_shard_h_copy(b200_groth16_pk_t handle, void *dev_buf, size_t first, size_t count, int to_key);
int b200_groth16_shard_h(b200_groth16_pk_t handle, void *dA, const void *dB, const void *dC);
int b200_groth16_shard_msm(b200_groth16_pk_t handle, unsigned mask, void *dev_blob);
/* `h` on three chain results and `msm` in one call: the pointwise step, the icosetFFT and the H shard MSM run beside
 * the other selected MSMs instead of after them. */
int b200_groth16_shard_finish_msm(b200_groth16_pk_t handle, void *dA, const void *dB, const void *dC, unsigned mask,
                                  void *dev_blob);
/* `msm` for the MSMs that need only the witness (mask over A, B.g2, B.g1, L; not H), started in the background: the call
 * returns at once, the caller runs the chain / broadcast / `h` / H-MSM steps of the same proof beside them (the reference
 * overlaps nothing here: r1cs_gg_ppzksnark.tcc:489-581 is sequential), and _end waits for the blob slots and reports the
 * first failure. One begin per key at a time; the key must not be freed in between. */
int b200_groth16_shard_msm_begin(b200_groth16_pk_t handle, unsigned mask, void *dev_blob);
int b200_groth16_shard_msm_end(b200_groth16_pk_t handle);
/* Optional, at the start of a sharded proof: starts the scalar multiplications of the assembly that need only r, s and
 * delta (r delta_g1, s delta_g1, s delta_g2, (r s) delta_g1; r1cs_gg_ppzksnark.tcc:583-595) on host threads, so that they
 * run beside the GPU work instead of after it. A later b200_groth16_assemble with the same curve, delta, r and s uses the
 * result; without a matching preparation it computes them itself. The inputs are copied. */
int b200_groth16_assemble_prepare(int curve, const void *delta_g1, const void *delta_g2, const void *r, const void *s);
/* blobs = `world` blobs in rank order (host memory). proof_out as b200_groth16_prove. */
int b200_groth16_assemble(int curve, const void *alpha_g1, const void *beta_g1, const void *beta_g2,
                          const void *delta_g1, const void *delta_g2, const size_t query_sizes[4], int world,
                          const void *blobs, const void *r, const void *s, void *proof_out);

/* ---- key generation: libff::batch_exp / batch_exp_with_coeff (multiexp.hpp, multiexp.tcc:879-946:
 *      res[i] = windowed_exp(table of g, coeff * v[i])) followed by libff::batch_to_special (:949-974), which is how
 *      r1cs_gg_ppzksnark_generator builds A_query, H_query, L_query and — through libsnark::kc_batch_exp
 *      (kc_multiexp.tcc:147-239) — B_query (r1cs_gg_ppzksnark.tcc:321-398).
 *      g: one point, reference layout, special form. coeff: one Fr (Montgomery) or NULL for 1. scalars: n Fr.
 *      out: n points, reference layout, special form: (x, y, one), or zero() = (0, one, 0) where coeff * v[i] = 0. */
int b200_batch_exp(int curve, int group, const void *g, const void *coeff, const void *scalars, size_t n,
                   void *out);

/* ---- measurement helpers (used by bench.py; not part of the reference surface) ------- */
/* Integer-pipe microbenchmarks, multiply-adds per second: wide = 32x32+64 IMAD.WIDE.U32(.X) chains
 * (what a Montgomery product is made of; two issue slots each), lo = 32-bit IMAD (the pipe's
 * issue rate). Operands are loop-variant so that nothing is hoisted. */
int b200_imad_peak(int iters, double *wide_mads_per_s, double *lo_mads_per_s);
/* Time (ms, CUDA events on the library stream) of the last MSM's kernels:
 * [0] scalar prep+sort, [1] bucket accumulation, [2] bucket reduction, [3] fold, [4] total */
int b200_last_msm_timing(double *ms5);
/* CUDA-event stopwatch on the library's stream (device time of everything launched between). */
int b200_timer_start(void);
int b200_timer_stop(double *ms);
/* Synthetic bases for benchmarks: out_host[i] = [k0 + i] * generator, i < n (k0 >= 1), affine,
 * reference layout; computed on the device. */
int b200_make_bases(int curve, int group, uint64_t k0, size_t n, void *out_host);
/* Wide multiply-adds per second of carry-chained IMAD.WIDE.X (valid after b200_imad_peak). */
double b200_imad_carry_peak(void);
/* Number of kernels this library has launched since b200_init (bench.py's gpu_launches). */
uint64_t b200_launch_count(void);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* LIBSNARK_B200_H_ */
