/*
 * merkle_oracle.c -- CPU ORACLE (test infrastructure, NOT product code): the "Merkle tree built" span of Commit.
 *
 * Restates core/tree.go of ChainSafe/lumenos (reference @ 9f33d62): calculateNodeHash (tree.go:37-74), NewTree /
 * NewTreeWithHashStrategy (:76-163), GetMerklePath (:176-219), with crypto/sha256 restated from FIPS 180-4, and the leaf
 * byte stream of processLeafParallel (fhe/ligero.go:157-158: ct.WriteTo(buf)) as
 *     header || 2 x ( poly_header || limbs x ( limb_header || N little-endian u64 ) ).
 * PARITY STATUS: SHA-256 is pinned by the FIPS 180-4 known answers and hashlib (tests/test_oracle.py); the tree shape by the
 * reference's own VerifyMerklePath rule (tree.go:223-268).  The framing bytes of lattigo's Ciphertext.WriteTo are NOT pinned
 * (lattigo is not vendored): they are a caller-supplied parameter here and in the product.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static const uint32_t K256[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5, 0xd807aa98, 0x12835b01, 0x243185be,
    0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174, 0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa,
    0x5cb0a9dc, 0x76f988da, 0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967, 0x27b70a85,
    0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85, 0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3,
    0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070, 0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f,
    0x682e6ff3, 0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};

static uint32_t ror(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }

static void sha256_block(uint32_t st[8], const uint8_t *p) {
    uint32_t w[64];
    for (int i = 0; i < 16; i++) w[i] = ((uint32_t)p[4 * i] << 24) | ((uint32_t)p[4 * i + 1] << 16) | ((uint32_t)p[4 * i + 2] << 8) | p[4 * i + 3];
    for (int i = 16; i < 64; i++) {
        uint32_t s0 = ror(w[i - 15], 7) ^ ror(w[i - 15], 18) ^ (w[i - 15] >> 3);
        uint32_t s1 = ror(w[i - 2], 17) ^ ror(w[i - 2], 19) ^ (w[i - 2] >> 10);
        w[i] = w[i - 16] + s0 + w[i - 7] + s1;
    }
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
    for (int i = 0; i < 64; i++) {
        uint32_t t1 = h + (ror(e, 6) ^ ror(e, 11) ^ ror(e, 25)) + ((e & f) ^ (~e & g)) + K256[i] + w[i];
        uint32_t t2 = (ror(a, 2) ^ ror(a, 13) ^ ror(a, 22)) + ((a & b) ^ (a & c) ^ (b & c));
        h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

/* crypto/sha256: Sum256 */
void orc_sha256(const uint8_t *data, uint64_t len, uint8_t out[32]) {
    uint32_t st[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
    uint64_t i = 0;
    for (; i + 64 <= len; i += 64) sha256_block(st, data + i);
    uint8_t tail[128];
    uint64_t r = len - i;
    memset(tail, 0, sizeof tail);
    memcpy(tail, data + i, r);
    tail[r] = 0x80;
    uint64_t tl = r + 9 <= 64 ? 64 : 128, bits = len * 8;
    for (int j = 0; j < 8; j++) tail[tl - 1 - j] = (uint8_t)(bits >> (8 * j));
    for (uint64_t j = 0; j < tl; j += 64) sha256_block(st, tail + j);
    for (int j = 0; j < 8; j++) { out[4 * j] = st[j] >> 24; out[4 * j + 1] = st[j] >> 16; out[4 * j + 2] = st[j] >> 8; out[4 * j + 3] = st[j]; }
}

/* the leaf buffer of processLeafParallel (ligero.go:157-158) for a ciphertext ct[2][nl][N]; returns its length (out may be NULL) */
uint64_t orc_leaf_bytes(const uint64_t *ct, int nl, int N, const uint8_t *hdr, uint32_t hdr_len, const uint8_t *ph, uint32_t ph_len,
                        const uint8_t *lh, uint32_t lh_len, uint8_t *out) {
    uint64_t n = 0;
    if (out) memcpy(out, hdr, hdr_len);
    n += hdr_len;
    for (int p = 0; p < 2; p++) {
        if (out) memcpy(out + n, ph, ph_len);
        n += ph_len;
        for (int l = 0; l < nl; l++) {
            if (out) memcpy(out + n, lh, lh_len);
            n += lh_len;
            const uint64_t *src = ct + ((size_t)p * nl + l) * N;
            if (out) for (int j = 0; j < N; j++) for (int b = 0; b < 8; b++) out[n + 8 * (size_t)j + b] = (uint8_t)(src[j] >> (8 * b));
            n += 8 * (uint64_t)N;
        }
    }
    return n;
}

/* NewTreeWithHashStrategy (tree.go:81-163) on the n leaf hashes: nodes = level 0 (the leaf hashes), then every parent level
   (an unpaired last node is hashed with itself, :122-130), the root last.  Returns the node count (nodes may be NULL). */
uint32_t orc_merkle_tree(const uint8_t *leaf_hashes, uint32_t n, uint8_t *nodes) {
    uint32_t total = 0, cnt = n, off = 0;
    if (n == 0) return 0;
    if (nodes) memcpy(nodes, leaf_hashes, (size_t)n * 32);
    total = n;
    while (cnt > 1) {
        uint32_t np = 0;
        for (uint32_t i = 0; i < cnt; i += 2) {
            uint32_t l = i, r = i + 1 < cnt ? i + 1 : i;
            if (nodes) {
                uint8_t comb[64];
                memcpy(comb, nodes + (size_t)(off + l) * 32, 32);
                memcpy(comb + 32, nodes + (size_t)(off + r) * 32, 32);
                orc_sha256(comb, 64, nodes + (size_t)(total + np) * 32);
            }
            np++;
        }
        off = total; total += np; cnt = np;
    }
    return total;
}

/* GetMerklePath (tree.go:176-219): sibling hashes from the leaf level up; returns the number of hashes */
uint32_t orc_merkle_path(const uint8_t *nodes, uint32_t n, uint32_t index, uint8_t *out) {
    uint32_t off = 0, cnt = n, i = index, d = 0;
    while (cnt > 1) {
        uint32_t sib = (i & 1) ? i - 1 : (i + 1 < cnt ? i + 1 : i);
        memcpy(out + 32 * (size_t)d, nodes + (size_t)(off + sib) * 32, 32);
        off += cnt; cnt = (cnt + 1) / 2; i >>= 1; d++;
    }
    return d;
}

/* EncryptedProof.WriteTo (fhe/ligero.go:660-704) with LigeroMetadata.WriteTo (:757-763) and MerklePath.WriteTo (core/tree.go:270-277):
   metadata || MatR || MatZ || QueriedCols || paths || root.  mat_r / mat_z: cols ciphertexts [2][nl_m][N_m]; queried: nq ciphertexts
   [2][nl_q][N_q]; framing (hdr, ph, lh) per group as in orc_leaf_bytes.  Returns the length (out may be NULL). */
uint64_t orc_proof_bytes(uint32_t rows, uint32_t cols, uint32_t rho_inv, uint32_t queries,
                         const uint64_t *mat_r, const uint64_t *mat_z, uint32_t n_mat, int nl_m, int N_m,
                         const uint8_t *mh, uint32_t mh_len, const uint8_t *mph, uint32_t mph_len, const uint8_t *mlh, uint32_t mlh_len,
                         const uint64_t *queried, uint32_t n_q, int nl_q, int N_q,
                         const uint8_t *qh, uint32_t qh_len, const uint8_t *qph, uint32_t qph_len, const uint8_t *qlh, uint32_t qlh_len,
                         const uint8_t *paths, uint64_t paths_len, const uint8_t *root, uint32_t root_len, uint8_t *out) {
    uint64_t n = 0;
    if (out) {
        /* binary.Write(buf, LittleEndian, uint32(Rows) / uint32(Cols) / uint8(RhoInv) / uint16(Queries)) */
        for (int i = 0; i < 4; i++) { out[i] = (uint8_t)(rows >> (8 * i)); out[4 + i] = (uint8_t)(cols >> (8 * i)); }
        out[8] = (uint8_t)rho_inv;
        out[9] = (uint8_t)queries; out[10] = (uint8_t)(queries >> 8);
    }
    n = 11;
    const uint64_t *mats[2] = {mat_r, mat_z};
    for (int g = 0; g < 2; g++)
        for (uint32_t i = 0; i < n_mat; i++)
            n += orc_leaf_bytes(mats[g] + (size_t)i * 2 * nl_m * N_m, nl_m, N_m, mh, mh_len, mph, mph_len, mlh, mlh_len, out ? out + n : NULL);
    for (uint32_t i = 0; i < n_q; i++)
        n += orc_leaf_bytes(queried + (size_t)i * 2 * nl_q * N_q, nl_q, N_q, qh, qh_len, qph, qph_len, qlh, qlh_len, out ? out + n : NULL);
    if (out && paths_len) memcpy(out + n, paths, paths_len);
    n += paths_len;
    if (out && root_len) memcpy(out + n, root, root_len);
    n += root_len;
    return n;
}
