/*
 * lumenos_b200.h -- C ABI of liblumenos_b200.so: the B200 (sm_100a) implementation of lumenos's
 * server-side homomorphic Ligero prover hot path.  Plain pointers and sizes only; every entry point
 * names the reference interface (file:line in ChainSafe/lumenos @ 9f33d62) it replaces.  The Go host
 * keeps keys, encryption, transcript, Merkle tree and wire format, and binds these symbols through cgo
 * (see INTEGRATION.md for the shim).
 *
 * Data conventions (identical to rlwe.Ciphertext.Value[p].Coeffs[limb][j], core/transcript.go:34-35):
 *   - residues are uint64, NTT domain, NON-Montgomery, canonical [0, q_limb);
 *   - a ciphertext at level L is the contiguous block  [poly 0..1][limb 0..L][coeff 0..N-1];
 *   - evaluation keys are NTT + MONTGOMERY form (Lattigo's in-memory form), laid out
 *     [digit 0..beta-1][component 0..1][limb 0..k+np-1 (Q limbs then P limbs)][coeff].
 *
 * Error convention: every function returns 0 on success, a negative lmn_status otherwise, and never
 * aborts the process (the reference escalates Go errors to panic, fhe/ligero.go:509-512; its cgo
 * precedent returns int status, vdec/prover.go:220-232).  lmn_last_error() gives the message of the last
 * failure on the calling thread.
 *
 * Threading: a context may be used from several host threads for DISTINCT stores; calls on one context
 * are serialised internally (the reference runs R and Z inner products from two goroutines,
 * fhe/ligero.go:231-242 -- use lmn_inner_product2 to fuse them instead).
 *
 * Lifetime of host buffers: work is queued on the context's stream and most calls return before it has run.
 * Small host inputs (plaintexts, index lists, scalars, framing bytes) are copied into library-owned pinned
 * staging before the call returns, so the caller may reuse or free them at once, pinned or not.  The BULK
 * host buffers of lmn_encode_host (host_matrix) and lmn_inner_product2_host (host_r, host_b) are read /
 * written by asynchronous copies and must stay valid and untouched until lmn_ctx_synchronize returns.
 * lmn_store_upload / download and the key loaders synchronise before returning.
 */
#ifndef LUMENOS_B200_H
#define LUMENOS_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct lmn_ctx lmn_ctx;     /* device context: moduli, NTT tables, keys, workspace  */
typedef struct lmn_store lmn_store; /* device-resident array of ciphertexts (ref: LigeroProver{Matrix,EncodedMatrix}, fhe/ligero.go:32-37) */
typedef struct lmn_plan lmn_plan;   /* compiled butterfly plan of fhe.NTT for one size (ref: fhe/ntt.go:20-281) */

typedef enum lmn_status {
    LMN_OK = 0,
    LMN_ERR_INVALID = -1,     /* bad argument / shape / level */
    LMN_ERR_CUDA = -2,        /* CUDA runtime failure (message has the cudaError string) */
    LMN_ERR_NOKEY = -3,       /* Galois / ring-switch key not loaded */
    LMN_ERR_UNSUPPORTED = -4, /* parameter outside what the kernels are instantiated for */
    LMN_ERR_NOMEM = -5
} lmn_status;

/* BGV parameters as the host holds them (ref: fhe/bfv.go:121-188 GenerateBGVParamsForNTT -> bgv.Parameters).
 * psi_q / psi_p are the primitive 2N-th roots of unity the host ring uses, in plain (non-Montgomery) form:
 * ring.ModExp(subRing.PrimitiveRoot, (q-1)/subRing.NthRoot, q).  The library derives its NTT tables from
 * them so that its NTT domain is the host's NTT domain. */
typedef struct lmn_params {
    uint32_t log_n;          /* ring degree N = 2^log_n: 6, 8, 10, 12, 13 or 14 (others: LMN_ERR_UNSUPPORTED) */
    uint32_t num_q;          /* k: number of Q primes (max level = k-1)                  */
    uint32_t num_p;          /* number of special primes P (reference: 2)                */
    const uint64_t *q;       /* [num_q] params.Q()                                       */
    const uint64_t *p;       /* [num_p] params.P()                                       */
    const uint64_t *psi_q;   /* [num_q]                                                  */
    const uint64_t *psi_p;   /* [num_p]                                                  */
    uint64_t t;              /* plaintext modulus (cmd/server/main.go:22)                */
    /* hypothesis switches for un-vendored Lattigo behaviours; 0xffffffff = library default */
    uint32_t scalar_centered;       /* default 1: Evaluator.Mul(ct,uint64) centres the scalar mod t      */
    uint32_t moddown_round_nearest; /* default 1: ModDown is a rounded division by P                      */
    uint32_t innersum_rowswap_last; /* default 1: rows==N -> InnerSum(N/2) then row swap                  */
} lmn_params;

/* ---------------- context ---------------- */
/* replaces fhe.NewBackendBFV (fhe/bfv.go:22-27): binds parameters to one GPU. device < 0 = current device */
int lmn_ctx_create(const lmn_params *params, int device, lmn_ctx **out);
void lmn_ctx_destroy(lmn_ctx *ctx);
const char *lmn_last_error(void);
/* run all work of this context on an existing CUDA stream (cudaStream_t); NULL = the context's own (non-blocking) stream;
 * pass cudaStreamLegacy (0x1) / cudaStreamPerThread (0x2) to name CUDA's default streams */
int lmn_ctx_set_stream(lmn_ctx *ctx, void *cuda_stream);
int lmn_ctx_synchronize(lmn_ctx *ctx);
/* number of kernels launched by this context since creation (bench.py's gpu_launches) */
uint64_t lmn_ctx_launch_count(const lmn_ctx *ctx);
/* columns processed together per inner-product / rescale batch (workspace = batch * ~(5k+4) limbs).  The client-side entry points
 * (lmn_witness_columns, lmn_encrypt_columns, lmn_decrypt_decode) do not follow it: they always work 256 columns at a time, i.e.
 * pooled workspaces of 256 * N words for values and plaintexts and 256 * (level+1) * N words for decryption, kept by the context. */
int lmn_ctx_set_batch(lmn_ctx *ctx, uint32_t jobs_per_batch);

/* No reference counterpart (Go's GC plays this role there).  The context pools its scratch between calls so that a steady-state
 * prover never allocates: batch workspaces of the inner products and rescales (batch * ~(5k+4) limbs), the serialised leaves of the
 * last Merkle build (n_leaves * leaf bytes: 4 GiB for 8192 leaves of 512 KiB), exchange staging of the communicator (2 chunks of
 * batch ciphertexts + one chunk of whole ciphertexts) and the client-side batch buffers.  A host that commits once and then keeps
 * the context for a long time calls this to hand that memory back; it waits for the context's queued work, frees the pools and
 * reports the bytes released (released_bytes may be NULL).  The pools regrow on the next call that needs them. */
int lmn_ctx_trim(lmn_ctx *ctx, size_t *released_bytes);

/* ---------------- device-resident ciphertext store ---------------- */
/* replaces []*rlwe.Ciphertext held between Commit and Prove (fhe/ligero.go:118-123) */
int lmn_store_create(lmn_ctx *ctx, uint32_t n_ct, uint32_t level, lmn_store **out);
/* wrap caller-owned device memory (n_ct * 2*(level+1)*N uint64), e.g. a torch tensor */
int lmn_store_wrap(lmn_ctx *ctx, void *device_ptr, uint32_t n_ct, uint32_t level, lmn_store **out);
void lmn_store_destroy(lmn_store *s);
uint32_t lmn_store_count(const lmn_store *s);
uint32_t lmn_store_level(const lmn_store *s);
void *lmn_store_device_ptr(const lmn_store *s);
size_t lmn_store_bytes(const lmn_store *s);
/* bulk copies of whole ciphertexts in the contiguous layout above; host memory may be pageable or pinned */
int lmn_store_upload(lmn_store *s, uint32_t first, uint32_t count, const uint64_t *host);
int lmn_store_download(const lmn_store *s, uint32_t first, uint32_t count, uint64_t *host);
/* per-limb copies: Go's Coeffs is [][]uint64 (one slice per limb), and cgo may not pass pointer-to-pointer */
int lmn_store_upload_limb(lmn_store *s, uint32_t ct, uint32_t poly, uint32_t limb, const uint64_t *host);
int lmn_store_download_limb(const lmn_store *s, uint32_t ct, uint32_t poly, uint32_t limb, uint64_t *host);
/* The same for MANY limbs in one call - the form a Lattigo host uses for whole stores (rlwe.Ciphertext.Value[p].Coeffs[limb] is one
 * slice per limb: core/utils.go:12-33, vdec/prover.go:110-149).  limb_ptrs is a table in C memory with count * 2 * (level+1)
 * entries in store order, limb_ptrs[((ct - first) * 2 + poly) * (level+1) + limb] -> N uint64 (pageable or pinned; a Go caller pins
 * the slices with runtime.Pinner for the duration of the call).  The limbs are packed into / unpacked from two
 * alternating 32 MiB pinned buffers by up to 8 host threads while the other buffer's one cudaMemcpyAsync runs on the context's copy
 * stream: no synchronisation per limb.  upload: the caller's memory is free when the call returns, the data is visible to work queued
 * on the context afterwards.  download: waits for the work already queued on the context and returns with the data in place. */
int lmn_store_upload_limbs(lmn_store *s, uint32_t first, uint32_t count, const uint64_t *const *limb_ptrs);
int lmn_store_download_limbs(const lmn_store *s, uint32_t first, uint32_t count, uint64_t *const *limb_ptrs);
/* pinned host staging owned by the library (cudaHostAlloc) */
int lmn_host_alloc(size_t bytes, void **out);
void lmn_host_free(void *p);

/* ---------------- Encode eval (fhe/code.go + fhe/ntt.go) ---------------- */
/* compile the butterfly/twiddle/permutation plan of nttInner(size = cols*rho_inv) over the plaintext field the
 * backend was built with: roots_forward[i] = backend.Field().RootForwardUint64(i) (core/field.go:45-47), field_n =
 * backend.Field().N() entries.  The reference builds the field as core.NewPrimeField(t, cols*2) (cmd/server/main.go:57,
 * fhe/ligero_test.go:58) and nttInner reads `step = Field().N()/size` and `idx %= Field().N()` from it (fhe/ntt.go:249-266),
 * so field_n is a parameter of its own, equal to cols*rho_inv only for rho_inv == 2; every table index the walk produces is
 * checked against field_n. */
int lmn_plan_create(lmn_ctx *ctx, uint32_t cols, uint32_t rho_inv, const uint64_t *roots_forward, uint32_t field_n, lmn_plan **out);
void lmn_plan_destroy(lmn_plan *plan);
/* ciphertext-level op counts the plan stands for (ServerBFV.MulCounter analogue, fhe/bfv.go:20,34-46) */
int lmn_plan_counts(const lmn_plan *plan, uint64_t *adds, uint64_t *subs, uint64_t *muls);
uint32_t lmn_plan_size(const lmn_plan *plan);
/* replaces fhe.Encode(matrix, rows, rhoInv, backend) (fhe/code.go:8-34): matrix has `cols` max-level cts;
 * zero_ct is the host-produced encryption of the zero vector (code.go:15-22, random => must come from the
 * host); encoded receives cols*rho_inv max-level cts in the order fhe.NTT returns them. */
int lmn_encode(lmn_ctx *ctx, const lmn_plan *plan, const lmn_store *matrix, const lmn_store *zero_ct,
               uint32_t zero_idx, lmn_store *encoded);
/* Encode span of Commit with the matrix still in (pinned) host memory, laid out as lmn_store_upload expects: the upload into
 * `matrix` (kept for Prove) runs limb by limb on the copy stream while the limbs that have landed are encoded */
int lmn_encode_host(lmn_ctx *ctx, const lmn_plan *plan, const uint64_t *host_matrix, lmn_store *matrix, const lmn_store *zero, uint32_t zero_idx, lmn_store *encoded);
/* replaces fhe.NTT(values, size, backend) (fhe/ntt.go:12-17) for len(values) == size == plan size:
 * out[i] = i-th ciphertext of the returned slice. in and out must be distinct stores. */
int lmn_ntt(lmn_ctx *ctx, const lmn_plan *plan, const lmn_store *in, lmn_store *out);

/* multi-GPU form of lmn_encode: every lane (poly, limb, coefficient) is an independent transform,
 * so a rank may hold a coefficient slice of every ciphertext.  Buffers are raw device pointers laid out
 * [ciphertext][poly 0..1][limb 0..k-1][2^log_coeffs coefficients]; matrix has plan-cols entries, zero one entry,
 * encoded cols*rho_inv entries. */
int lmn_encode_lanes(lmn_ctx *ctx, const lmn_plan *plan, const void *matrix_dev, const void *zero_dev, void *encoded_dev,
                     uint32_t log_coeffs);

/* host-only self-check of the plan compiler (no device): evaluates the compiled plan of size n on a plaintext
 * vector over Z_t in place, which must reproduce core.NTT (core/ntt.go:3-6) -- the equality the reference's own
 * TestEncode asserts (fhe/code_test.go:100-116).  counts3 (optional) receives {adds, subs, muls}. */
int lmn_plan_host_check(uint32_t n, const uint64_t *roots_forward, uint64_t t, uint64_t *values, uint64_t *counts3);

/* ---------------- keys ---------------- */
/* replaces rlwe.MemEvaluationKeySet lookup by Galois element (fhe/ligero_test.go:53-56,
 * cmd/server/main.go:66-122): evk = gk.Value[d][0][c] Q then P limbs, Montgomery + NTT form */
int lmn_load_galois_key(lmn_ctx *ctx, uint64_t gal_el, const uint64_t *evk_host);
/* ring-switch key (fhe/ring_switch.go:16-57, 98-104): evk sk -> embed(sk_small), laid out [rns_digits][base2_digits][2][k+np][N]
 * = evk.Value[d][b][c].  The reference asks for BaseTwoDecomposition = 13 (ring_switch.go:47-55); whether Lattigo honours it
 * when P is present is unpinned, so the digit shape is an argument: rns_digits must equal ceil(k/np) and
 * base2_digits must be 1 (LMN_ERR_UNSUPPORTED otherwise - a key with base-2 sub-digits is rejected instead of being misread).
 * psi_small = primitive 2N'-th root of q[0] used by the small ring (0 = derive as psi_q[0]^(N/N')) */
int lmn_load_ring_switch_key(lmn_ctx *ctx, uint32_t log_n_small, const uint64_t *evk_host, uint32_t rns_digits, uint32_t base2_digits,
                             uint64_t psi_small);

/* ---------------- Inner product eval (fhe/ligero.go:299-370 matrixInnerSumEval) ---------------- */
/* pt_host: plaintext at max level as bgv.Encoder.Encode produced it: [num_q][N], NTT domain (ligero.go:204-205).
 * For each column i: MulNew(matrix[i], pt) -> InnerSum(.,1,rows) -> Rescale down to level 1 (ligero.go:319-333).
 * out must hold matrix-count level-1 ciphertexts.  The host multiplies its Scale metadata by
 * prod_{l=2..L} q_l^-1 mod t (returned by lmn_rescale_scale). */
int lmn_inner_product(lmn_ctx *ctx, const lmn_store *matrix, const uint64_t *pt_host, uint32_t rows, lmn_store *out);
/* fused pair: InnerProduct(Matrix, r) and InnerProduct(Matrix, b) reading the matrix once (ligero.go:231-242) */
int lmn_inner_product2(lmn_ctx *ctx, const lmn_store *matrix, const uint64_t *pt_r_host, const uint64_t *pt_b_host,
                       uint32_t rows, lmn_store *out_r, lmn_store *out_b);
/* the same with the results also copied to caller-owned (pinned, lmn_host_alloc) host buffers batch by batch on the context's copy
 * stream, behind the kernels of the following batches; complete when the context's stream is synchronised */
int lmn_inner_product2_host(lmn_ctx *ctx, const lmn_store *matrix, const uint64_t *pt_r, const uint64_t *pt_b, uint32_t rows,
                            lmn_store *out_r, lmn_store *out_b, uint64_t *host_r, uint64_t *host_b);
/* optional ring switch of every level-1 ciphertext of `in` (ligero.go:336-342, ring_switch.go:106-113):
 * out is a level-0 store created with lmn_store_create_small */
int lmn_store_create_small(lmn_ctx *ctx, uint32_t n_ct, lmn_store **out); /* [2][1][N'] per ct */
int lmn_ring_switch(lmn_ctx *ctx, const lmn_store *in_level1, lmn_store *out_small);

/* ---------------- Query columns eval (fhe/ligero.go:262-280) ---------------- */
/* out[i] = EncodedMatrix[idx[i]] rescaled to level 1 (duplicates allowed; the reference's in-place second
 * rescale of a duplicate is a no-op, so duplicates yield identical ciphertexts). */
int lmn_query_columns(lmn_ctx *ctx, const lmn_store *encoded, const uint32_t *idx, uint32_t n_idx, lmn_store *out);
/* product of q_l^-1 mod t for l = from_level .. to_level+1: the factor by which bgv Rescale updates Scale */
uint64_t lmn_rescale_scale(const lmn_ctx *ctx, uint32_t from_level, uint32_t to_level);

/* ---------------- Commit eval: leaves + Merkle tree (fhe/ligero.go:105-116, 126-183; core/tree.go:37-219) ---------------- */
/* leaves[i] = encoded[i] rescaled to level 1: the CopyNew + Rescale loop of processLeafParallel (ligero.go:146-154).
 * The leaves store also serves the query stage: lmn_store_gather(leaves, idx) == lmn_query_columns(encoded, idx). */
int lmn_commit_leaves(lmn_ctx *ctx, const lmn_store *encoded, lmn_store *leaves);
/* out[i] = src[idx[i]] (same level) */
int lmn_store_gather(lmn_ctx *ctx, const lmn_store *src, const uint32_t *idx, uint32_t n_idx, lmn_store *out);
/* Byte stream of one leaf = what ct.WriteTo(buf) emits (ligero.go:157-158):
 *   header || for each of the 2 polynomials: poly_header || for each limb: limb_header || N coefficients as little-endian u64.
 * The framing bytes are the caller's: lattigo's (MetaData bytes, slice-length prefixes) are produced by the Go host once per
 * Commit (they do not depend on the column); poly_header_len, limb_header_len <= 16. */
typedef struct {
    const uint8_t *header; uint32_t header_len;
    const uint8_t *poly_header; uint32_t poly_header_len;
    const uint8_t *limb_header; uint32_t limb_header_len;
} lmn_leaf_format;
typedef struct lmn_merkle lmn_merkle;
/* core.NewTree(leafs) (tree.go:76-163): SHA-256 leaf hashes, parents = SHA-256(left || right), an unpaired last node is paired
 * with itself.  Hashing runs on the GPU; the node hashes (32 B each) are kept on the host for root and paths. */
int lmn_merkle_build(lmn_ctx *ctx, const lmn_store *leaves, const lmn_leaf_format *fmt, lmn_merkle **out);
/* multi-GPU Commit (columns sharded by rank): each rank hashes the leaves of its own columns (32 bytes per leaf, to host memory),
 * the hashes are gathered in column order and the tree is built from them; lmn_merkle_build == the two calls on one GPU */
int lmn_leaf_hashes(lmn_ctx *ctx, const lmn_store *leaves, const lmn_leaf_format *fmt, uint8_t *out);
int lmn_merkle_from_leaf_hashes(lmn_ctx *ctx, const uint8_t *hashes, uint32_t n_leaves, lmn_merkle **out);
/* No counterpart in the reference: the same tree as lmn_merkle_from_leaf_hashes for hosts that must not allocate while they run -
 * a prover that has committed all of its HBM up front (stores, workspaces, another pool beside this library), so that a cudaMalloc
 * in the middle of Commit may fail, or one whose other threads have kernels in flight, which every cudaMalloc / cudaFree would
 * stall.  The call makes no cudaMalloc / cudaFree / new: the tree's nodes are built in a device array that is part of the library
 * image (2 MiB per GPU, held until the process exits) and the results are written into the caller's host buffers:
 *   root_out32          the 32-byte root (== lmn_merkle_root);
 *   nodes_out, optional every node, 32 bytes each, level by level from the leaf hashes up - level l holds ceil(n_leaves / 2^l)
 *                       nodes and the last level is the root - so that paths can be read off as lmn_merkle_path does (the sibling
 *                       of node i is i ^ 1, or i itself when that is past the end of the level); nodes_cap = its size in bytes.
 * n_leaves above LMN_MERKLE_NOALLOC_MAX_LEAVES is rejected with LMN_ERR_INVALID.  The first call on each GPU also resolves the
 * array's device address; make it during set-up, before any stream capture begins.  The array is shared by all contexts of the
 * process and keeps the last call's nodes (every entry a call reads is written by that call first); calls are serialised by a
 * process-wide lock and return when their results are in host memory, because two that overlapped on the GPU would race. */
#define LMN_MERKLE_NOALLOC_MAX_LEAVES 32768u
int lmn_merkle_from_leaf_hashes_noalloc(lmn_ctx *ctx, const uint8_t *hashes, uint32_t n_leaves, uint8_t *root_out32, uint8_t *nodes_out,
                                        size_t nodes_cap);
void lmn_merkle_destroy(lmn_merkle *tree);
uint32_t lmn_merkle_leaf_count(const lmn_merkle *tree);
uint32_t lmn_merkle_depth(const lmn_merkle *tree);                            /* hashes per path */
int lmn_merkle_root(const lmn_merkle *tree, uint8_t *out32);                  /* MerkleTree.MerkleRoot, tree.go:165-172 */
int lmn_merkle_leaf_hash(const lmn_merkle *tree, uint32_t index, uint8_t *out32);
int lmn_merkle_path(const lmn_merkle *tree, uint32_t index, uint8_t *out);    /* GetMerklePath, tree.go:176-219: 32*depth bytes */
/* ---------------- Wire format of the proof (fhe/ligero.go:646-704, 757-763; core/tree.go:270-277) ---------------- */
/* ct.WriteTo(s[first]) || ... || ct.WriteTo(s[first+count-1]) written to host memory `out` (framing as for the leaves; any host
 * memory, pinned memory overlaps the copy with the serialisation).  out may be NULL to query *len.  Multi-GPU hosts marshal each
 * rank's columns into its own range of one buffer (every ciphertext stream has the same length). */
int lmn_store_marshal(lmn_ctx *ctx, const lmn_store *s, uint32_t first, uint32_t count, const lmn_leaf_format *fmt, uint8_t *out, size_t cap, size_t *len);
typedef struct { uint32_t rows, cols, rho_inv, queries; } lmn_ligero_metadata;   /* fhe.LigeroMetadata, ligero.go:65-70 */
/* EncryptedProof.WriteTo / MarshalBinary: metadata (u32 rows, u32 cols, u8 rhoInv, u16 queries, little-endian) || MatR || MatZ ||
 * QueriedCols (each ciphertext as ct.WriteTo) || paths (MerklePath.WriteTo: the 32-byte hashes back to back, `paths_len` bytes as
 * returned by lmn_merkle_path per query) || root.  fmt_query == NULL: same framing as MatR / MatZ (they differ after a ring switch).
 * out may be NULL to query *len. */
int lmn_proof_marshal(lmn_ctx *ctx, const lmn_ligero_metadata *md, const lmn_store *mat_r, const lmn_store *mat_z, const lmn_store *queried,
                      const lmn_leaf_format *fmt_mat, const lmn_leaf_format *fmt_query, const uint8_t *paths, size_t paths_len,
                      const uint8_t *root, size_t root_len, uint8_t *out, size_t cap, size_t *len);
/* test hook: the serialised bytes of leaf `index` (out may be NULL to query *len) */
int lmn_leaf_bytes(lmn_ctx *ctx, const lmn_store *leaves, uint32_t index, const lmn_leaf_format *fmt, uint8_t *out, size_t cap, size_t *len);

/* ---------------- ServerBFV operator level (fhe/bfv.go:13-58; promoted bgv.Evaluator methods) ---------------- */
/* each operand is (store, index); stores must be at the same level unless noted */
int lmn_ct_add(lmn_ctx *ctx, const lmn_store *a, uint32_t ai, const lmn_store *b, uint32_t bi, lmn_store *o, uint32_t oi); /* Evaluator.Add, ntt.go:27 */
int lmn_ct_sub(lmn_ctx *ctx, const lmn_store *a, uint32_t ai, const lmn_store *b, uint32_t bi, lmn_store *o, uint32_t oi); /* Evaluator.Sub, ntt.go:31 */
int lmn_ct_mul_scalar(lmn_ctx *ctx, const lmn_store *a, uint32_t ai, uint64_t w, lmn_store *o, uint32_t oi);              /* ServerBFV.Mul(ct,uint64), bfv.go:34-37 */
int lmn_ct_mul_pt(lmn_ctx *ctx, const lmn_store *a, uint32_t ai, const uint64_t *pt_host, lmn_store *o, uint32_t oi);      /* ServerBFV.MulNew(ct,pt), bfv.go:39-42 */
int lmn_ct_inner_sum(lmn_ctx *ctx, lmn_store *a, uint32_t ai, uint32_t n);                                                 /* Evaluator.InnerSum(ct,1,n,ct), ligero.go:325 */
int lmn_ct_rescale(lmn_ctx *ctx, const lmn_store *a, uint32_t ai, lmn_store *o, uint32_t oi);                              /* Evaluator.Rescale: o is one level below a, ligero.go:332 */
int lmn_ct_copy(lmn_ctx *ctx, const lmn_store *a, uint32_t ai, lmn_store *o, uint32_t oi);                                 /* Ciphertext.CopyNew, code.go:12 */
/* ring.SubRing NTT / INTT of one limb of one polynomial, in place (test hook for the limb kernels) */
int lmn_limb_ntt(lmn_ctx *ctx, lmn_store *a, uint32_t ai, uint32_t poly, uint32_t limb, int inverse);

/* ---------------- multi-GPU exchange (one process per GPU, NCCL over NVLink) ---------------- */
/* The inner products, the query rescale and the leaf rescale shard by COLUMN BLOCK (rank r owns ciphertexts [r n/G, (r+1) n/G));
 * Encode shards by COEFFICIENT SLICE (rank r holds coefficients [r N/G, (r+1) N/G) of every limb of every ciphertext, laid out
 * [ciphertext][poly][limb][N/G], see lmn_encode_lanes).  The calls below are the layout changes between the two, as grouped
 * ncclSend / ncclRecv all-to-alls on the communicator's own stream, ordered with the context's stream by events.  Every rank must
 * make the same calls in the same order with the same batch size (lmn_ctx_set_batch bounds the chunk of an exchange).  NCCL is
 * resolved at run time (dlopen of libnccl.so.2: the copy the process already has loaded, else the system's). */
int lmn_comm_unique_id(uint8_t *out128);                                    /* ncclGetUniqueId on one rank; the host distributes it */
int lmn_comm_init(lmn_ctx *ctx, const uint8_t *id128, int rank, int world); /* ncclCommInitRank on the context's device */
int lmn_comm_destroy(lmn_ctx *ctx);
int lmn_comm_rank(const lmn_ctx *ctx);
int lmn_comm_world(const lmn_ctx *ctx);
/* Prove without a leaves store (fhe/ligero.go:262-280): whole max-level ciphertexts of the queried columns this rank rescales
 * (query positions rank, rank + G, ... of query_idx, which is identical on all ranks) from every rank's slice of the encoded
 * matrix (n ciphertexts); out holds ceil((n_queries - rank) / G) entries in list order */
int lmn_exchange_queried(lmn_ctx *ctx, const void *enc_slice_dev, uint32_t n, const uint32_t *query_idx, uint32_t n_queries, lmn_store *out);
/* the whole encoded matrix: slices -> this rank's column block (n / G max-level ciphertexts) */
int lmn_exchange_encoded(lmn_ctx *ctx, const void *enc_slice_dev, uint32_t n, lmn_store *block);
/* Commit's leaves (fhe/ligero.go:126-183) of this rank's column block straight from the slices: the all-to-all of one chunk runs
 * while the previous chunk is rescaled to level 1; == lmn_exchange_encoded + lmn_commit_leaves without the intermediate store */
int lmn_commit_leaves_from_slices(lmn_ctx *ctx, const void *enc_slice_dev, uint32_t n, lmn_store *leaves);
/* the inverse: this rank's column block of n / G ciphertexts -> its coefficient slice of all n ciphertexts (a matrix uploaded or
 * encrypted once as column blocks feeds the slice-wise Encode) */
int lmn_exchange_to_slices(lmn_ctx *ctx, const lmn_store *block, uint32_t n, void *slice_dev);
/* Merkle leaf hashes of all ranks in column order: mine = 32 n_local bytes (lmn_leaf_hashes), all = 32 n_local G bytes (host) */
int lmn_allgather_leaf_hashes(lmn_ctx *ctx, const uint8_t *mine, uint32_t n_local, uint8_t *all);

/* ---------------- the stages either side of the path ---------------- */
/* Keys and the plaintext ring arrive as Lattigo holds them: psi_t = the primitive 2N-th root of unity of the encoder's ring T
 * (ring.SubRing of t, as psi_q above); pk = pk.Value[0].Q, pk.Value[1].Q as [2][num_q][N] and sk = sk.Value.Q as [num_q][N],
 * both NTT + MONTGOMERY form (rlwe.PublicKey / rlwe.SecretKey in memory). */
int lmn_load_plaintext_ring(lmn_ctx *ctx, uint64_t psi_t);
int lmn_load_public_key(lmn_ctx *ctx, const uint64_t *pk_host);
int lmn_load_secret_key(lmn_ctx *ctx, const uint64_t *sk_host);
/* core.RandomMatrixRowMajor's witness (core/utils.go:46-82) generated on the device: ChaCha20 (IETF; key = chacha_key32, zero nonce,
 * counter 0) keystream as little-endian u64 words mod t, row-major rows x cols; host_out receives columns [first_col,
 * first_col + n_cols) column-major, [n_cols][rows].  The reference's key is LE64(1) followed by 24 zero bytes. */
int lmn_witness_columns(lmn_ctx *ctx, const uint8_t *chacha_key32, uint32_t rows, uint32_t cols, uint32_t first_col, uint32_t n_cols,
                        uint64_t *host_out);
/* "Encrypt matrix" span of the demo server (cmd/server/main.go:187-207) on the device.  For j in [0, count):
 *   values = column first_col + j of the witness (values_colmajor_host [count][rows] when given, any u64, reduced mod t like
 *            ringT.Reduce; else generated on the device from chacha_key32 as lmn_witness_columns does, rows x cols_total);
 *   matrix[j] = EncryptNew(Encode(values)) at max level: ct = (u pk0 + e0 + pt, u pk1 + e1), u uniform ternary, e0 / e1 rounded
 *            Gaussians (sigma 3.2, bound 19 = lattigo DefaultXe).
 * All randomness comes from the ChaCha20 stream keyed by enc_seed32 with nonce (column index, domain tag | stream, attempt), so a
 * column's ciphertext does not depend on which rank or batch produced it.  Lattigo's own sampler stream cannot be reproduced (not
 * vendored): the contract is that the client decrypts every column to the witness, with Lattigo-sized noise - not ciphertext bits.
 * SECURITY: enc_seed32 is the ONLY source of encryption randomness.  It must be 32 bytes of fresh CSPRNG output (crypto/rand) for
 * every matrix and must never be used again: two matrices encrypted under one seed give ct_j - ct'_j = (pt_j - pt'_j, 0), which
 * reveals the difference of the plaintexts without the key (the reference's EncryptNew draws fresh randomness on every call).  All
 * ranks that encrypt column blocks of ONE matrix share its seed; the zero column of that matrix (lmn_encrypt_zero) may share it too:
 * the nonce carries a domain tag (column / zero column), and the witness keystream uses the all-zero nonce, so the three uses never
 * read the same ChaCha20 block even if enc_seed32 and chacha_key32 are the same bytes. */
int lmn_encrypt_columns(lmn_ctx *ctx, const uint64_t *values_colmajor_host, const uint8_t *chacha_key32, uint32_t rows, uint32_t cols_total,
                        uint32_t first_col, uint32_t count, const uint8_t *enc_seed32, lmn_store *matrix);
/* zeroCol of fhe.Encode (fhe/code.go:15-22): EncryptNew of the all-zero plaintext into out[out_idx].  stream_id plays the column
 * index in the nonce, under the zero-column domain tag: any value is safe against the matrix columns; two zero columns made from one
 * seed need distinct stream_ids.  The seed rule of lmn_encrypt_columns applies (fresh per matrix, never reused). */
int lmn_encrypt_zero(lmn_ctx *ctx, const uint8_t *enc_seed32, uint32_t stream_id, lmn_store *out, uint32_t out_idx);
/* test hook: the three small polynomials the encryptor draws for one ciphertext - u (uniform ternary), e0, e1 (rounded Gaussians) -
 * as signed coefficients, out3_host [3][N].  id = the column index (zero_column = 0) or lmn_encrypt_zero's stream_id (zero_column = 1).
 * The encryption of zero made from the same seed and id satisfies c0 + c1 s = u e_pk + e0 + e1 s exactly, which is how the tests tie
 * this dump to the ciphertexts.  Never call it with a seed that encrypts real data: u, e0 and e1 open the ciphertext. */
int lmn_test_encryption_samples(lmn_ctx *ctx, const uint8_t *enc_seed32, uint32_t id, int zero_column, int64_t *out3_host);
/* backend.Encode(values, pt) at max level (fhe/ligero.go:204-205, 218-219: the r and b plaintexts of Prove): pt_host [num_q][N] */
int lmn_encode_plaintext(lmn_ctx *ctx, const uint64_t *values, uint32_t n_values, uint64_t *pt_host);
/* client side: EncryptedProof.Decrypt's decryptBatchedParallel (fhe/ligero.go:381-502, 577-636) = rlwe.Decryptor.DecryptNew +
 * bgv.Encoder.Decode of ciphertexts [first, first + count) of a store of the context ring (any level); values_host receives the
 * first n_values slots of each, [count][n_values].  scale = the ciphertexts' rlwe.MetaData.Scale (mod t). */
int lmn_decrypt_decode(lmn_ctx *ctx, const lmn_store *s, uint32_t first, uint32_t count, uint64_t scale, uint32_t n_values,
                       uint64_t *values_host);

/* ---------------- measurement hooks (bench.py) ---------------- */
/* per-kernel-class device timing with CUDA event pairs on the launch stream; mask bit i = class i */
int lmn_profile_enable(lmn_ctx *ctx, uint32_t mask);
/* ms: summed device time of the timed launches of the class; launches / jobs: launches since the last reset and the column-jobs
 * (ciphertexts) they processed (block-pipeline classes; 0 for the others) */
int lmn_profile_read(lmn_ctx *ctx, int kernel_id, double *ms, uint64_t *launches, uint64_t *jobs, int reset);
int lmn_kernel_count(void);
const char *lmn_kernel_name(int kernel_id);
/* measured ceiling of the NTT butterfly on the integer pipes: the fastest formulation found on this part (3-product Shoup quotient,
 * middle sum folded into IMAD.WIDE's 64-bit addend, plain adds, lazy [0, 3q); tools/int_pipe_probe.cu k_bfly<1>), register-resident,
 * 64 warps per SM, nothing else in the loop */
int lmn_bench_int_peak(lmn_ctx *ctx, double *butterflies_per_s);
/* the same network with the butterfly the block kernels are built with (ntt_blk.cuh bfly_ct3), for comparison with the ceiling */
int lmn_bench_product_butterfly(lmn_ctx *ctx, double *butterflies_per_s);

#ifdef __cplusplus
}
#endif
#endif
