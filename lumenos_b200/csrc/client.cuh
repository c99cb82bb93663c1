// Launchers of kernels_client.cu: the two stages either side of the server hot path.
//   row 3  "Encrypt matrix" of the demo server (cmd/server/main.go:187-207): core.RandomMatrixRowMajor's ChaCha20 witness,
//          bgv.Encoder.Encode of every column and rlwe.Encryptor.EncryptNew under the public key, all on the device;
//   row 4  the client's EncryptedProof.Decrypt (fhe/ligero.go:381-502, 577-636): rlwe.Decryptor + bgv.Encoder.Decode.
#pragma once
#include "kernels.cuh"

namespace lmn {

struct ChaChaKey { uint32_t w[8]; };            // 32-byte key as little-endian words

// plaintext ring Z_t[X]/(X^N + 1) of the bgv.Encoder (lattigo: ringT) and the slot permutation
struct PlainRing {
    ModDev mod;                                  // t
    const ulonglong2 *tw_fwd, *tw_inv;           // [N] (w, floor(w 2^64 / t)), Lattigo table order
    const uint32_t *perm;                        // [N] bgv slot i -> coefficient-domain position before the INTT
};

constexpr int CRT_MAX = 28;
struct CrtConsts {                               // Garner mixed-radix reconstruction of x mod Q_level, then mod t (bgv RingQ2T)
    int nl;
    u64 half[CRT_MAX];                           // (q_i - 1) / 2
    u64 inv_w[CRT_MAX][CRT_MAX], inv_ws[CRT_MAX][CRT_MAX];   // [i][j] q_j^-1 mod q_i, j < i (Shoup pair)
    u64 q_mod_t[CRT_MAX];                        // q_i mod t
    u64 half_q_mod_t;                            // (Q - 1)/2 mod t
    u64 t_w[CRT_MAX], t_ws[CRT_MAX];             // N^-1 t mod q_i (Shoup pair): INTT scaling and the BGV "x t" in one product
    u64 sinv_w, sinv_ws;                         // Scale^-1 mod t
};

// W[j][i] = witness(row i, column first_col + j) = LE64 word (i cols + first_col + j) of the ChaCha20 keystream, mod t
void launch_witness(const Launch &L, const ChaChaKey &key, u64 t, uint32_t rows, uint32_t cols, uint32_t first_col, uint32_t n_cols, u64 *W);
// m[j][c] = coefficient c of the plaintext polynomial of values[j][0..n_values): perm scatter, INTT over Z_t, times N^-1 (canonical)
void launch_pt_encode(const Launch &L, int logn, const PlainRing &R, const u64 *values, uint32_t n_values, uint64_t value_stride, uint32_t n_cols, u64 *m);
// pt[j][i][.] = NTT_qi(m[j] t^-1 mod q_i): the NTT-domain plaintext at `nl` limbs (bgv.Encoder.Encode output)
void launch_pt_lift(const Launch &L, const DeviceTables &T, int nl, const u64 *m, uint32_t n_cols, const ulonglong2 *tinv, u64 *pt);
// ct[j] = (u pk0 + e0 + pt_j, u pk1 + e1), u ternary, e0 / e1 Gaussian (sigma 3.2, bound 19) drawn from the ChaCha20 stream of
// `seed` with nonce (column id first_id + j, domain | stream, attempt); m may be null (encryption of zero).  pk: [2][k][N] NTT +
// Montgomery.  `domain` separates the users of one seed: a matrix column, the zero column and the witness keystream (nonce 0, 0, 0)
// never share a ChaCha20 block, whatever ids the caller picks.
constexpr uint32_t ENC_DOMAIN_COLUMN = 0x100u, ENC_DOMAIN_ZERO = 0x200u;
void launch_encrypt(const Launch &L, const DeviceTables &T, const ChaChaKey &seed, uint32_t first_id, uint32_t domain, uint32_t n_cols, const u64 *m,
                    const ulonglong2 *tinv, const u64 *pk, u64 *cts, uint64_t ct_words);
// test hook: out3[0..2][N] = the signed coefficients of u, e0, e1 that launch_encrypt draws for column id `id` under `domain`
void launch_dump_small(const Launch &L, const DeviceTables &T, const ChaChaKey &seed, uint32_t id, uint32_t domain, u64 q, long long *out3);
// buf[j][i][.] = INTT_qi(c0 + c1 s) N^-1 t : the coefficient-domain t-multiple of the decryption (s: [k][N] NTT + Montgomery)
void launch_decrypt_limbs(const Launch &L, const DeviceTables &T, const CrtConsts *Cd, int nl, const u64 *cts, uint64_t ct_words, uint32_t n_ct,
                          const u64 *sk, u64 *buf);
// values[j][0..n_values) = slots of round(buf_j mod Q -> t) Scale^-1: Garner, centre, NTT over Z_t, un-permute
void launch_crt_decode(const Launch &L, int logn, const PlainRing &R, const CrtConsts *Cd, const ModDev *mods, const u64 *buf, uint32_t n_ct,
                       uint32_t n_values, u64 *values);
const char *kernels_client_init();

}  // namespace lmn
