/*
 * lumenos_oracle.h -- CPU ORACLE (test infrastructure, NOT product code).
 *
 * Plain-C restatement of the server-side homomorphic Ligero prover hot path of
 * ChainSafe/lumenos (reference @ 9f33d62) and of the Lattigo v6 ring/rlwe/bgv
 * arithmetic it calls (github.com/tuneinsight/lattigo/v6
 * v6.1.2-0.20250520151126-84f6bc33cb5b, go.mod:5 -- NOT vendored in the reference,
 * restated from its published algorithms).
 *
 * PARITY STATUS: pinned against the reference's own known answers for the witness
 * generator (P(1) values, results/baseline/client/bench_*.txt:22) and, at decrypted-
 * plaintext level, against the reference's own plaintext twin (core/ntt.go, core/code.go,
 * fhe/ligero.go:799-953 LigeroProveReference) exactly as the reference's tests do
 * (fhe/code_test.go:109-116, fhe/ligero_test.go:164-174).  Ciphertext-bit parity against
 * real Lattigo is "parity unpinned": the reference holds no golden ciphertexts (keys and
 * encryption noise are fresh-random in every reference test) and neither Go nor Lattigo's
 * source exist in this environment (probed on the GPU box too, profiles/r2_go_probe.txt).
 * Structural facts of the Lattigo half that reference-held data does pin (tests/test_host_logic.py,
 * fixture tests/golden/logged_sizes.json): the byte length of Ciphertext.WriteTo's framing (the logged
 * sizes of MatR / MatZ / QueriedCols / proof), the evaluation-key shape [ceil(k/2)][2][k+2][N] without
 * base-2 sub-digits, one such key for the ring switch, and the Galois set of InnerSum including the row
 * swap at rows == N (the logged key-blob sizes).  What stays open is listed in DESIGN.md section 5.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this library.  The product (lumenos_b200) never links or calls it.
 */
#ifndef LUMENOS_ORACLE_H
#define LUMENOS_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_params orc_params;

/* hypothesis switches for behaviours of un-vendored Lattigo */
typedef struct orc_switches {
    int primroot_start;        /* first candidate g of ring.PrimitiveRoot: 3 (default) or 2   */
    int scalar_centered;       /* bgv Evaluator.Mul(ct,uint64): centre scalar mod t (default 1) */
    int digit_centered;        /* multi-prime key-switch digit centred (default 0)              */
    int moddown_round_nearest; /* ModDown adds P/2 before conversion (default 1)                */
    int innersum_rowswap_last; /* rows==N: InnerSum(N/2) then row swap (1) or swap first (0)    */
} orc_switches;

void orc_default_switches(orc_switches *sw);

/* ---- parameters (fhe/bfv.go:121-188 + Lattigo rlwe.GenModuli) ---- */
/* k primes LogQ=[58,56,...], 2 primes LogP=[55,55] generated Lattigo-style */
orc_params *orc_params_generate(int logN, int k, uint64_t t, const orc_switches *sw);
/* explicit moduli (drop-in mode: host passes params.Q(), params.P()) */
orc_params *orc_params_from_moduli(int logN, int k, const uint64_t *q, int np, const uint64_t *p,
                                   uint64_t t, const orc_switches *sw);
void orc_params_free(orc_params *P);
int orc_params_k(const orc_params *P);
int orc_params_np(const orc_params *P);
int orc_params_N(const orc_params *P);
uint64_t orc_params_t(const orc_params *P);
void orc_params_moduli(const orc_params *P, uint64_t *q, uint64_t *p);
void orc_params_psi(const orc_params *P, uint64_t *psi_q, uint64_t *psi_p); /* plain (non-Montgomery) 2N-th roots */
uint64_t orc_params_psi_t(const orc_params *P);                             /* the same for the plaintext ring Z_t[X]/(X^N+1) */
/* fhe.GenerateBGVParamsForNTT heuristic: number of Q primes for a given nttSize (=cols) */
int orc_num_q_primes_for_ntt(int nttSize, uint64_t t);

/* plaintext prover's row combinations (core/math.go:8-21, fhe/ligero.go:799-953): out[j] = sum_i r_i M[i][j] mod t */
void orc_vec_mat_mod(const uint64_t *r, const uint64_t *M_rowmajor, int rows, int cols, uint64_t t, uint64_t *out);

/* ---- witness generator (core/utils.go:46-82) ---- */
void orc_witness_matrix(int rows, int cols, uint64_t t, uint64_t *out_rowmajor);

/* ---- plaintext field twin (core/field.go, core/ntt.go, core/code.go) ---- */
/* RootsForward table of core.NewPrimeField(t, Nfield): Nfield entries, Montgomery form */
int orc_field_roots_forward(uint64_t t, int Nfield, int primroot_start, uint64_t *table);
/* core.NTT(values,size) in place, len(values)=len, field built with N=Nfield */
void orc_core_ntt(uint64_t *v, int len, int size, uint64_t t, const uint64_t *roots, int Nfield);
/* core.Encode(row, rhoInv): out has cols*rhoInv entries */
void orc_core_encode(const uint64_t *row, int cols, int rhoInv, uint64_t t, const uint64_t *roots,
                     uint64_t *out);

/* ---- limb-level ring ops (Lattigo ring.SubRing) ---- */
/* mod_idx: 0..k-1 = Q primes, k..k+np-1 = P primes */
void orc_ntt_limb(const orc_params *P, int mod_idx, uint64_t *a);
void orc_intt_limb(const orc_params *P, int mod_idx, uint64_t *a);
void orc_automorphism_index(const orc_params *P, uint64_t galEl, uint32_t *index);

/* ---- BGV client side (test infrastructure only: keys, encrypt, decrypt) ---- */
/* sk: N int8 ternary coefficients */
void orc_keygen_sk(const orc_params *P, uint64_t seed, int8_t *sk);
/* pk: [2][k][N] NTT, non-Montgomery, over Q only */
void orc_keygen_pk(const orc_params *P, const int8_t *sk, uint64_t seed, uint64_t *pk);
/* Galois key for galEl: [beta][2][k+np][N], NTT + Montgomery form (Lattigo layout evk.Value[d][0][c]) */
int orc_beta(const orc_params *P);
void orc_keygen_galois(const orc_params *P, const int8_t *sk, uint64_t galEl, uint64_t seed,
                       uint64_t *evk);
/* evaluation key sk -> sk_small (ring switch, fhe/ring_switch.go:16-57): sk_small has Nsmall coeffs */
void orc_keygen_ringswitch(const orc_params *P, const int8_t *sk, const int8_t *sk_small, int logNsmall,
                           uint64_t seed, uint64_t *evk);
/* bgv.Encoder.Encode at level `level`: values[n] -> pt[(level+1)][N] (NTT, scaled by T^-1, scale=1) */
void orc_encode_pt(const orc_params *P, const uint64_t *values, int n, int level, uint64_t *pt);
/* rlwe.Encryptor (pk) at max level: ct [2][k][N]; pt may be NULL (encryption of zero) */
void orc_encrypt_pk(const orc_params *P, const uint64_t *pk, const uint64_t *pt, uint64_t seed,
                    uint64_t *ct);
/* decrypt + bgv decode: ct [2][level+1][N] at `level`, scale = tracked bgv scale; values[N] */
void orc_decrypt_decode(const orc_params *P, const int8_t *sk, const uint64_t *ct, int level,
                        uint64_t scale, uint64_t *values);
/* noise budget helper: max over coefficients of log2 |T*(c0+c1 s) - scale*m| is not needed by tests;
   returns log2 of the centred infinity norm of (c0 + c1*s) * T mod Q_level minus message -- diagnostic */
double orc_noise_bits(const orc_params *P, const int8_t *sk, const uint64_t *ct, int level);

/* ---- ServerBFV operator level (fhe/bfv.go + Lattigo bgv.Evaluator) ---- */
/* all cts are [2][level+1][N] contiguous, NTT domain, canonical */
void orc_ct_add(const orc_params *P, int level, const uint64_t *a, const uint64_t *b, uint64_t *out);
void orc_ct_sub(const orc_params *P, int level, const uint64_t *a, const uint64_t *b, uint64_t *out);
void orc_ct_mul_scalar(const orc_params *P, int level, const uint64_t *a, uint64_t w, uint64_t *out);
void orc_ct_mul_pt(const orc_params *P, int level, const uint64_t *a, const uint64_t *pt, uint64_t *out);
/* InnerSum(ct,1,n) in place; evks[i] = galois key for galEls[i] */
int orc_galois_elements_for_innersum(const orc_params *P, int n, uint64_t *galEls);
int orc_ct_innersum(const orc_params *P, int level, uint64_t *ct, int n, const uint64_t *galEls,
                    const uint64_t *const *evks, int nkeys);
/* Rescale by last modulus: in [2][level+1][N] -> out [2][level][N]; returns new scale */
uint64_t orc_ct_rescale(const orc_params *P, int level, const uint64_t *in, uint64_t *out, uint64_t scale);
/* single key-switched automorphism (rlwe.Evaluator.Automorphism) */
void orc_ct_automorphism(const orc_params *P, int level, const uint64_t *in, uint64_t galEl,
                         const uint64_t *evk, uint64_t *out);
/* ApplyEvaluationKey into a smaller ring (fhe/ring_switch.go:106-113): in level-1 ct, out level-0 ct of
   degree 2^logNsmall: [2][1][Nsmall] */
void orc_ct_ringswitch(const orc_params *P, const uint64_t *in_lvl1, const uint64_t *evk, int logNsmall,
                       uint64_t *out);

/* ---- path level (fhe/ntt.go, fhe/code.go, fhe/ligero.go) ---- */
/* fhe.NTT(values,size): cts[n][2][k][N] contiguous in/out (pointer permutations applied) */
void orc_fhe_ntt(const orc_params *P, uint64_t *cts, int n, int size, const uint64_t *roots, int Nfield);
/* fhe.Encode: matrix[cols][2][k][N], zero ct -> encoded[cols*rhoInv][2][k][N] */
void orc_fhe_encode(const orc_params *P, const uint64_t *matrix, int cols, int rhoInv,
                    const uint64_t *zero_ct, const uint64_t *roots, uint64_t *encoded);
/* restricted to a lane range [lane0,lane1) of the 2*k*N lanes (for bounded CPU baselines / big shapes):
   matrix_lanes[cols][nl], zero_lanes[nl] -> encoded_lanes[cols*rhoInv][nl]; limb_of_lane[nl] gives the Q index */
void orc_fhe_encode_lanes(const orc_params *P, const uint64_t *matrix_lanes, int cols, int rhoInv,
                          const uint64_t *zero_lanes, const int *limb_of_lane, int nl,
                          const uint64_t *roots, uint64_t *encoded_lanes);
/* matrixInnerSumEval for one column: ct[2][k][N] x pt[k][N] -> out level-1 ct [2][2][N]; returns scale */
uint64_t orc_inner_sum_eval_column(const orc_params *P, const uint64_t *ct, const uint64_t *pt, int rows,
                                   const uint64_t *galEls, const uint64_t *const *evks, int nkeys,
                                   uint64_t *out);
/* rescale max-level ct down to level 1 (query columns / leaves); returns scale */
uint64_t orc_rescale_to_level1(const orc_params *P, const uint64_t *ct, uint64_t *out);

/* op counters of the last orc_fhe_ntt call (add, sub, mul) -- SURVEY App. D totals */
void orc_fhe_ntt_counts(uint64_t *adds, uint64_t *subs, uint64_t *muls);

/* ---- "Merkle tree built" span of Commit (merkle_oracle.c: core/tree.go, fhe/ligero.go:126-183) ---- */
void orc_sha256(const uint8_t *data, uint64_t len, uint8_t out[32]);
uint64_t orc_leaf_bytes(const uint64_t *ct, int nl, int N, const uint8_t *hdr, uint32_t hdr_len, const uint8_t *ph, uint32_t ph_len,
                        const uint8_t *lh, uint32_t lh_len, uint8_t *out);
uint32_t orc_merkle_tree(const uint8_t *leaf_hashes, uint32_t n, uint8_t *nodes);
uint32_t orc_merkle_path(const uint8_t *nodes, uint32_t n, uint32_t index, uint8_t *out);
/* EncryptedProof.WriteTo (fhe/ligero.go:660-704) */
uint64_t orc_proof_bytes(uint32_t rows, uint32_t cols, uint32_t rho_inv, uint32_t queries,
                         const uint64_t *mat_r, const uint64_t *mat_z, uint32_t n_mat, int nl_m, int N_m,
                         const uint8_t *mh, uint32_t mh_len, const uint8_t *mph, uint32_t mph_len, const uint8_t *mlh, uint32_t mlh_len,
                         const uint64_t *queried, uint32_t n_q, int nl_q, int N_q,
                         const uint8_t *qh, uint32_t qh_len, const uint8_t *qph, uint32_t qph_len, const uint8_t *qlh, uint32_t qlh_len,
                         const uint8_t *paths, uint64_t paths_len, const uint8_t *root, uint32_t root_len, uint8_t *out);

#ifdef __cplusplus
}
#endif
#endif
