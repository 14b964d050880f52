/*
 * gpufhe_b200 -- C ABI of the B200-native negacyclic-NTT / RNS polynomial-multiplication path.
 *
 * This is the drop-in boundary for GPU-FHE's polynomial-ring hot path.  The reference has no
 * FFI of its own (it is three pure-Python modules); each entry point below is the native
 * implementation behind one reference-level Python function, which keeps its signature in
 * arithmetic.py / primitive.py / polynomial.py at the repo root:
 *
 *   gfhe_vec_add / _sub / _mul   <- /root/reference/arithmetic.py:3-5, :7-9, :11-13
 *   gfhe_ntt_forward             <- /root/reference/arithmetic.py:15-16  (NTT, an identity stub there)
 *   gfhe_ntt_inverse             <- /root/reference/arithmetic.py:18-19  (iNTT, an identity stub there)
 *   gfhe_plan_create             <- /root/reference/primitive.py:3-4     (placeholder for the roots of unity)
 *   gfhe_poly_mul                <- /root/reference/" polynomial.py":3-5 (module that would hold poly_mul)
 *   gfhe_poly_mul_pre            <- the same product with one operand cached in the NTT domain
 *   gfhe_poly_mul_static         <- no counterpart in the reference: gfhe_poly_mul on a static scratch array, for callers
 *                                   that cannot allocate
 *   gfhe_ntt_mul_acc             <- /root/reference/arithmetic.py:11-13 applied in the NTT domain (sum of pointwise
 *                                   products; the reference has no NTT-domain API -- SURVEY.md section 8f row 2)
 *   gfhe_automorphism            <- absent from the reference: a(X) -> a(X^k)
 *   gfhe_*_host                  <- the same calls made with host (numpy) arrays, as the reference's
 *                                   callers would make them; copies are pipelined inside the call
 *   gfhe_vec_*_u64               <- /root/reference/arithmetic.py:3-13 on uint64 containers (values >= 2^63 are positive)
 *   gfhe_plan_export / _import   <- the table cache /root/reference/primitive.py:3-4 was the placeholder for, in a
 *                                   serialisable form
 *   gfhe_peer_*, gfhe_push_gather <- absent from the reference (no collective of any kind): the
 *                                   north-star's "final all-gather of limbs back to the caller's device" over NVLink
 *
 * Conventions: plain pointers and sizes only; all device pointers are raw CUDA device addresses
 * (e.g. torch.Tensor.data_ptr()); `stream` is a cudaStream_t passed as void* (NULL = legacy
 * default stream); work is enqueued asynchronously unless the name ends in _host.  Every function
 * returns 0 on success, a negative GFHE_E* code for argument errors, or a positive cudaError_t.
 * No exception crosses this boundary.  The caller owns every buffer; the library owns only plans.
 *
 * Data layout: residues are 64-bit words, canonical tensor layout [L, B, N] -- limb outermost,
 * N contiguous -- so one limb's batch is one contiguous block.  NTT(x)[k] = x(psi^(2*brv(k)+1)):
 * forward output is in bit-reversed evaluation order, inverse output is natural order in [0, q).
 */
#ifndef GPUFHE_B200_H
#define GPUFHE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GFHE_ABI_VERSION 6

#define GFHE_EARG      (-1)   /* null pointer / zero size */
#define GFHE_ELOGN     (-2)   /* log2(N) outside [4, 20] or missing roots */
#define GFHE_EMODULUS  (-3)   /* modulus outside [2, 2^62) */
#define GFHE_ENOTNTT   (-4)   /* modulus is not 1 mod 2N */
#define GFHE_EROOT     (-5)   /* psi^N != -1 mod q */
#define GFHE_ENOMEM    (-6)
#define GFHE_ERANGE    (-7)   /* limb range outside the plan */
#define GFHE_EWORKSPACE (-8)  /* workspace missing or too small */
#define GFHE_ENOTABLES (-9)   /* plan was created without NTT tables (log2_n == 0) */
#define GFHE_ETERMS    (-10)  /* too many terms: terms * q^2 must fit 128 bits */
#define GFHE_EGALOIS   (-11)  /* Galois element must be odd */
#define GFHE_ELAYOUT   (-12)  /* packed 32-bit layout asked of a plan with a modulus >= 2^30 */
#define GFHE_EALIGN    (-13)  /* a device buffer of a transform / product call is not 16-byte aligned */
#define GFHE_EFORMAT   (-14)  /* gfhe_plan_import: not a plan image of this library version, or corrupted */
#define GFHE_ESTATICWS (-15)  /* gfhe_poly_mul_static: the job needs more scratch than GFHE_STATIC_WS_BYTES */
#define GFHE_ENOALLOC  (-16)  /* gfhe_poly_mul_static: the plan runs the pipelined kernels (GFHE_PIPE=1), which allocate */

typedef struct gfhe_plan* gfhe_plan_t;   /* opaque; one per (device, N, modulus chain) */

int gfhe_abi_version(void);
const char* gfhe_error_string(int code);

/* Build per-limb constants and, if log2_n != 0, twiddle + Shoup tables (generated on the host
 * from psi[l], a primitive 2N-th root of unity mod moduli[l]) and upload them to `device`.
 * log2_n == 0 creates a modulus set usable by gfhe_vec_* only (psi may be NULL). */
int gfhe_plan_create(gfhe_plan_t* out, int device, uint32_t log2_n, uint32_t num_limbs,
                     const uint64_t* moduli, const uint64_t* psi);
int gfhe_plan_destroy(gfhe_plan_t plan);
int gfhe_plan_info(gfhe_plan_t plan, int* device, uint32_t* log2_n, uint32_t* num_limbs);
/* The plan's moduli and roots (psi[l] = 0 for a plan without tables): writes min(n, num_limbs) entries, returns the count. */
int gfhe_plan_moduli(gfhe_plan_t plan, uint64_t* moduli, uint64_t* psi, uint32_t n);

/* Serialisable table cache (the reference's primitive.py:3-4 has no tables at all).
 * gfhe_plan_export writes a self-describing image (header, moduli, roots, per-limb constants and the twiddle tables as
 * they sit in HBM, checksummed) into a caller buffer of gfhe_plan_export_bytes(); gfhe_plan_import validates an image
 * and uploads it to `device` -- no root search and no table generation.  GFHE_EFORMAT: not an image of this format,
 * truncated, or corrupted; GFHE_EWORKSPACE: export buffer too small. */
size_t gfhe_plan_export_bytes(gfhe_plan_t plan);
int gfhe_plan_export(gfhe_plan_t plan, void* buf, size_t bytes);
int gfhe_plan_import(gfhe_plan_t* out, int device, const void* buf, size_t bytes);

/* Batched transforms over limbs [limb0, limb0 + num_limbs) of the plan; buffers are
 * [num_limbs, batch, N].  in == out is allowed.  Inputs must be canonical residues in [0, q) (both directions: the
 * lazy-reduction range schedule starts from q); every device buffer of the transform and product calls must be
 * 16-byte aligned (GFHE_EALIGN otherwise -- the kernels use 16-byte vector accesses and TMA bulk copies). */
int gfhe_ntt_forward(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint32_t batch,
                     const uint64_t* d_in, uint64_t* d_out, void* stream);
int gfhe_ntt_inverse(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint32_t batch,
                     const uint64_t* d_in, uint64_t* d_out, void* stream);

/* c = a * b in Z_q[X]/(X^N + 1) per limb, coefficient domain in and out, as one fused pipeline
 * (pointwise product and N^-1 folded into the adjacent NTT stages).  c may alias neither a nor b.
 * N > 4096 needs a device workspace of gfhe_poly_mul_workspace_bytes(). */
size_t gfhe_poly_mul_workspace_bytes(gfhe_plan_t plan, uint32_t num_limbs, uint32_t batch);
int gfhe_poly_mul(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint32_t batch,
                  const uint64_t* d_a, const uint64_t* d_b, uint64_t* d_c,
                  void* d_workspace, size_t workspace_bytes, void* stream);

/* Same product when b is ALREADY in the NTT domain (d_b_ntt = what gfhe_ntt_forward wrote for b, canonical residues):
 * c = iNTT(NTT(a) .* d_b_ntt).  b's forward transform is skipped (a third of the work) -- for callers that multiply many
 * polynomials by one cached operand.  No workspace.  c may alias neither input. */
int gfhe_poly_mul_pre(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint32_t batch,
                      const uint64_t* d_a, const uint64_t* d_b_ntt, uint64_t* d_c, void* stream);

/* gfhe_poly_mul for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same product, same arguments minus the workspace: the operand-sized scratch of N > 4096 is a file-scope device array
 * inside the library, GFHE_STATIC_WS_BYTES (32 MiB) on every GPU that loads the library, held until the process exits.
 * A job needs num_limbs * batch * N * 8 bytes of it (N <= 4096: none); a larger job is refused with GFHE_ESTATICWS --
 * [4, 16, 65536], the config-4 shard of one of 8 GPUs, is exactly the limit.  The call makes no cudaMalloc / cudaFree and
 * creates no stream or event: three kernel launches on `stream` (no two-stream split); plans created under GFHE_PIPE=1
 * are refused with GFHE_ENOALLOC.  The first call on a device resolves the array's device address (make one call before
 * starting a stream capture).  The scratch keeps the last call's words and needs no clearing (the product overwrites
 * every word before reading it), but ALL calls on one GPU share it: two calls that overlap on the GPU (different streams,
 * different threads or plans) race -- order them on one stream or with events.  c may alias neither a nor b. */
#define GFHE_STATIC_WS_BYTES ((size_t)32 << 20)
size_t gfhe_static_workspace_bytes(void);
int gfhe_poly_mul_static(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint32_t batch,
                         const uint64_t* d_a, const uint64_t* d_b, uint64_t* d_c, void* stream);

/* Packed 32-bit layout: the same four operations on [num_limbs, batch, N] arrays of 32-bit words (int32 / uint32
 * tensors), for plans whose moduli are ALL below 2^30 (GFHE_ELAYOUT otherwise) -- the north-star's "32-bit where the
 * moduli fit".  Half the HBM traffic of the 64-bit layout; results are the same residues.  The poly-mul workspace is
 * gfhe_poly_mul_workspace_bytes() / 2. */
int gfhe_ntt_forward_u32(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint32_t batch,
                         const uint32_t* d_in, uint32_t* d_out, void* stream);
int gfhe_ntt_inverse_u32(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint32_t batch,
                         const uint32_t* d_in, uint32_t* d_out, void* stream);
int gfhe_poly_mul_u32(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint32_t batch,
                      const uint32_t* d_a, const uint32_t* d_b, uint32_t* d_c,
                      void* d_workspace, size_t workspace_bytes, void* stream);
int gfhe_poly_mul_pre_u32(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint32_t batch,
                          const uint32_t* d_a, const uint32_t* d_b_ntt, uint32_t* d_c, void* stream);

/* ... and the operations around them on the same layout: element-wise ops on ANY int32 inputs (Python % semantics,
 * canonical outputs), the NTT-domain multiply-accumulate (up to 65536 terms) and the automorphism.  Same contracts as
 * gfhe_vec_add/_sub/_mul, gfhe_ntt_mul_acc and gfhe_automorphism below. */
int gfhe_vec_add_i32(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint64_t inner,
                     const int32_t* d_a, const int32_t* d_b, int32_t* d_out, void* stream);
int gfhe_vec_sub_i32(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint64_t inner,
                     const int32_t* d_a, const int32_t* d_b, int32_t* d_out, void* stream);
int gfhe_vec_mul_i32(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint64_t inner,
                     const int32_t* d_a, const int32_t* d_b, int32_t* d_out, void* stream);
int gfhe_ntt_mul_acc_u32(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint64_t inner, uint32_t terms,
                         const uint32_t* d_a, const uint32_t* d_b, uint64_t term_stride, uint32_t* d_out, void* stream);
int gfhe_automorphism_u32(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint32_t batch, uint32_t k, int ntt_domain,
                          const uint32_t* d_in, uint32_t* d_out, void* stream);

/* out = (a op b) mod q_l, exact for ANY int64 inputs (Python % semantics, result in [0, q)).
 * Buffers are [num_limbs, inner]; limb l of the range uses modulus limb0 + l. */
int gfhe_vec_add(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint64_t inner,
                 const int64_t* d_a, const int64_t* d_b, int64_t* d_out, void* stream);
int gfhe_vec_sub(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint64_t inner,
                 const int64_t* d_a, const int64_t* d_b, int64_t* d_out, void* stream);
int gfhe_vec_mul(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint64_t inner,
                 const int64_t* d_a, const int64_t* d_b, int64_t* d_out, void* stream);

/* The same for arrays of UNSIGNED 64-bit words (uint64 containers): values >= 2^63 are large positive numbers here,
 * negative ones above.  Same kernels, the input reduction differs. */
int gfhe_vec_add_u64(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint64_t inner,
                     const uint64_t* d_a, const uint64_t* d_b, uint64_t* d_out, void* stream);
int gfhe_vec_sub_u64(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint64_t inner,
                     const uint64_t* d_a, const uint64_t* d_b, uint64_t* d_out, void* stream);
int gfhe_vec_mul_u64(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint64_t inner,
                     const uint64_t* d_a, const uint64_t* d_b, uint64_t* d_out, void* stream);

/* NTT-domain multiply-accumulate: out[l, i] = sum_{t < terms} a[t*term_stride + l*inner + i] * b[same] mod q_l.
 * a, b hold `terms` arrays of layout [num_limbs, inner] (inner = batch * N) spaced term_stride words apart; inputs
 * must be canonical residues in [0, q) (what gfhe_ntt_forward writes).  terms == 1 is the pointwise product.
 * Products are accumulated in 128 bits and reduced once: terms * q^2 must be < 2^128 (GFHE_ETERMS otherwise;
 * 256 terms for 60-bit moduli).  out may alias a or b only when terms == 1. */
int gfhe_ntt_mul_acc(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint64_t inner, uint32_t terms,
                     const uint64_t* d_a, const uint64_t* d_b, uint64_t term_stride, uint64_t* d_out, void* stream);

/* Galois automorphism a(X) -> a(X^k) of every limb-polynomial, k odd (taken mod 2N).  ntt_domain == 0: coefficient
 * domain in and out (signed index permutation); ntt_domain != 0: both in the bit-reversed evaluation order of
 * gfhe_ntt_forward (pure permutation), so that automorphism(NTT(a)) == NTT(automorphism(a)).  Out of place. */
int gfhe_automorphism(gfhe_plan_t plan, uint32_t limb0, uint32_t num_limbs, uint32_t batch, uint32_t k, int ntt_domain,
                      const uint64_t* d_in, uint64_t* d_out, void* stream);

/* The path's one collective, the all-gather of result limbs to every GPU of one box, as NVLink peer stores issued by the
 * library (gather.cu) instead of a separate NCCL call.  One process per GPU: each rank allocates its full [L, B, N] result
 * array with gfhe_peer_alloc, which also returns a 64-byte CUDA IPC handle; the ranks exchange the handles (any transport)
 * and open each other's arrays with gfhe_peer_open.  gfhe_push_gather then copies `bytes` from d_src (this rank's result
 * block, or one limb of it) to offset dst_offset_bytes of every array in dst_bases (host array of ndst <= GFHE_MAX_PEERS
 * device pointers: the peers', and this rank's own if it wants the block there too), asynchronously on `stream` -- so it
 * can follow each limb's kernels while the next limb is computed.  It never waits for another GPU; the caller orders
 * "all pushes have landed" with one stream-ordered collective of its own after the last push (shard.py, PeerGather).
 * `ctas` caps the copy kernel's grid: 0 = one CTA per SM (fastest alone), about 32 when the push overlaps other kernels.
 * Pointers, sizes and offsets must be multiples of 16 bytes (GFHE_EALIGN). */
#define GFHE_MAX_PEERS 16
int gfhe_peer_alloc(int device, size_t bytes, void** d_ptr, unsigned char* handle64);
int gfhe_peer_open(int device, const unsigned char* handle64, void** d_ptr);
int gfhe_peer_close(int device, void* d_ptr);
int gfhe_peer_free(int device, void* d_ptr);
int gfhe_push_gather(int device, const void* d_src, size_t bytes, void* const* dst_bases, uint32_t ndst,
                     size_t dst_offset_bytes, uint32_t ctas, void* stream);

/* Host-buffer variants: h_* are host arrays [L, batch, N] covering ALL limbs of the plan.  Limb groups are streamed
 * H2D -> kernels -> D2H on separate streams so copies overlap compute.  Pinned arrays are copied from / to in place;
 * pageable arrays (plain numpy / malloc memory) are staged through an internal ring of pinned buffers by a few host
 * threads (GFHE_HOST_COPY_THREADS, default half the cores, at most 8) while a drain thread copies finished groups out, so the three streams still overlap.  Synchronous: nothing is in flight
 * on the caller's buffers when the call returns, on error paths included. */
int gfhe_ntt_host(gfhe_plan_t plan, int inverse, uint32_t batch, const uint64_t* h_in, uint64_t* h_out);
int gfhe_poly_mul_host(gfhe_plan_t plan, uint32_t batch, const uint64_t* h_a, const uint64_t* h_b, uint64_t* h_c);

/* What the last gfhe_*_host call on this plan spent where: out[k] for k < min(n, GFHE_HOST_STATS_N), returns the count.
 * The *_MS entries of the copies and kernels are stream-busy times summed over limb groups (CUDA events around each
 * group's copies / launches), so bytes / ms is the rate each direction actually ran at. */
#define GFHE_HOST_STAT_H2D_MS       0
#define GFHE_HOST_STAT_KERNEL_MS    1
#define GFHE_HOST_STAT_D2H_MS       2
#define GFHE_HOST_STAT_H2D_BYTES    3
#define GFHE_HOST_STAT_D2H_BYTES    4
#define GFHE_HOST_STAT_WALL_MS      5
#define GFHE_HOST_STAT_STAGE_IN_MS  6   /* host threads copying pageable inputs into pinned staging */
#define GFHE_HOST_STAT_STAGE_OUT_MS 7   /* ... and results out of it */
#define GFHE_HOST_STAT_STAGED       8   /* 1 when the call staged (pageable caller), else 0 */
#define GFHE_HOST_STAT_GROUPS       9
#define GFHE_HOST_STATS_N           10
int gfhe_host_stats(gfhe_plan_t plan, double* out, uint32_t n);

/* Test hook (no GPU work): the ticket order of the pipelined kernels (ntt_pipe.inl) for `passes` passes with
 * tiles_per_pass[k] tiles per polynomial, npoly polynomials and `resident` resident CTAs.  Writes (pass, polynomial,
 * tile) triples for the first max_tickets tickets, the lag chosen (polynomials between a pass and its consumer), and
 * returns the ticket count. */
int gfhe_debug_pipe_schedule(uint32_t passes, const uint32_t* tiles_per_pass, uint32_t npoly, uint32_t resident,
                             uint32_t* out, uint32_t max_tickets, uint32_t* lag);

/* Measurement hook (bench.py): when enabled, every kernel launched through `plan` is bracketed
 * by cudaEvents on its stream.  gfhe_profile_read waits for them, writes up to max_kernels
 * (kernel name, summed milliseconds, launch count) triples and returns how many it wrote
 * (>= 0); a negative return is an error (GFHE_E*, or -cudaError_t). */
int gfhe_profile_enable(gfhe_plan_t plan, int on);
int gfhe_profile_read(gfhe_plan_t plan, uint32_t max_kernels, const char** names, double* total_ms,
                      uint64_t* launches);

#ifdef __cplusplus
}
#endif
#endif /* GPUFHE_B200_H */
