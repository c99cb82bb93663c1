// C++ twin of the reference's TestEncode / TestLigeroE2E (fhe/code_test.go:14-123, fhe/ligero_test.go:70-176) written against
// the C++ host mirror (include/lumenos_b200.hpp).  The CPU oracle plays Lattigo: it makes the keys, encrypts the witness,
// and provides the expected ciphertexts / decryptions.  Exit code 0 = all checks passed.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "lumenos_b200.hpp"
#include "lumenos_oracle.h"

using namespace lumenos;
typedef uint64_t u64;

#define REQUIRE(cond, ...) do { if (!(cond)) { std::fprintf(stderr, "FAIL %s:%d: ", __FILE__, __LINE__); std::fprintf(stderr, __VA_ARGS__); std::fprintf(stderr, "\n"); return 1; } } while (0)

static const u64 Modulus = 144115188075593729ull;  // fhe/ligero_test.go:16

int main(int argc, char **argv) {
    const int logN = argc > 1 ? std::atoi(argv[1]) : 8, rows = argc > 2 ? std::atoi(argv[2]) : 256, cols = argc > 3 ? std::atoi(argv[3]) : 16,
              k = argc > 4 ? std::atoi(argv[4]) : 6, rhoInv = 2;
    orc_switches sw; orc_default_switches(&sw);
    orc_params *P = orc_params_generate(logN, k, Modulus, &sw);
    REQUIRE(P, "oracle parameters");
    const int N = orc_params_N(P), np = orc_params_np(P), n = cols * rhoInv;
    const size_t ctw = 2 * (size_t)k * N;

    // ---- host side (Lattigo's job in the reference): keys, witness, encryption ----
    std::vector<int8_t> sk(N);
    orc_keygen_sk(P, 11, sk.data());
    std::vector<u64> pk(ctw);
    orc_keygen_pk(P, sk.data(), 12, pk.data());
    std::vector<u64> M((size_t)rows * cols);
    orc_witness_matrix(rows, cols, Modulus, M.data());
    std::vector<u64> cts(ctw * cols), zero(ctw), col(rows), pt((size_t)k * N);
    for (int j = 0; j < cols; j++) {
        for (int i = 0; i < rows; i++) col[i] = M[(size_t)i * cols + j];
        orc_encode_pt(P, col.data(), rows, k - 1, pt.data());
        orc_encrypt_pk(P, pk.data(), pt.data(), 100 + j, cts.data() + ctw * j);
    }
    orc_encrypt_pk(P, pk.data(), nullptr, 13, zero.data());
    u64 galEls[64];
    const int ngal = orc_galois_elements_for_innersum(P, rows, galEls);
    const int beta = orc_beta(P);
    std::vector<std::vector<u64>> evks(ngal, std::vector<u64>((size_t)beta * 2 * (k + np) * N));
    std::vector<const u64 *> evkp(ngal);
    for (int i = 0; i < ngal; i++) { orc_keygen_galois(P, sk.data(), galEls[i], 14, evks[i].data()); evkp[i] = evks[i].data(); }
    fhe::PrimeField field; field.modulus = Modulus; field.roots.resize(n);
    REQUIRE(orc_field_roots_forward(Modulus, n, sw.primroot_start, field.roots.data()) == 0, "field tables");
    std::vector<u64> r(rows), b(rows), rpt((size_t)k * N), bpt((size_t)k * N);
    u64 seed = 0x9e3779b97f4a7c15ull;
    for (int i = 0; i < rows; i++) { seed = seed * 6364136223846793005ull + 1442695040888963407ull; r[i] = seed; }   // unreduced, like SampleUints
    u64 zc = 1; for (int i = 0; i < cols; i++) zc = (u64)((unsigned __int128)zc * 3 % Modulus);                       // z = 3, z^cols
    u64 pw = 1; for (int i = 0; i < rows; i++) { b[i] = pw; pw = (u64)((unsigned __int128)pw * zc % Modulus); }
    orc_encode_pt(P, r.data(), rows, k - 1, rpt.data());
    orc_encode_pt(P, b.data(), rows, k - 1, bpt.data());

    // ---- the path under test, through the C++ mirror ----
    fhe::Parameters params;
    params.LogN = logN; params.T = Modulus;
    params.Q.resize(k); params.P.resize(np); params.PsiQ.resize(k); params.PsiP.resize(np);
    orc_params_moduli(P, params.Q.data(), params.P.data());
    orc_params_psi(P, params.PsiQ.data(), params.PsiP.data());
    params.PsiT = orc_params_psi_t(P);
    try {
        fhe::ServerBFV server(field, params, 0);
        for (int i = 0; i < ngal; i++) server.LoadGaloisKey(galEls[i], evks[i].data());
        fhe::CiphertextStore matrix(server, cols), zeroCol(server, 1);
        matrix.Upload(cts.data(), 0, cols);
        zeroCol.Upload(zero.data(), 0, 1);
        fhe::LigeroCommitter ligero(128, rows, cols, rhoInv);
        REQUIRE(ligero.Queries == 309, "queries = %d", ligero.Queries);
        fhe::LeafFormat lf;                         // lattigo-like framing: MetaData placeholder + u64 slice-length prefixes
        lf.header = {'m', 'e', 't', 'a', 0x01, 2, 0, 0, 0, 0, 0, 0, 0};
        lf.polyHeader.assign(8, 0); lf.polyHeader[0] = 2;
        lf.limbHeader.assign(8, 0); for (int b = 0; b < 4; b++) lf.limbHeader[b] = (uint8_t)((u64)N >> (8 * b));
        fhe::LigeroProver comm = ligero.Commit(matrix, server, {&zeroCol, 0}, &lf);
        std::vector<uint32_t> idx(ligero.Queries);
        for (int i = 0; i < ligero.Queries; i++) { seed = seed * 6364136223846793005ull + 1442695040888963407ull; idx[i] = (uint32_t)((seed >> 33) % n); }
        idx[9] = idx[4];
        fhe::EncryptedProof proof = comm.Prove(rpt.data(), bpt.data(), idx, server);

        // TestEncode: decrypt(Encode) == core.Encode, and bit-exact against the oracle's fhe.Encode
        std::vector<u64> enc_want(ctw * n), enc_got = comm.EncodedMatrix->Download();
        orc_fhe_encode(P, cts.data(), cols, rhoInv, zero.data(), field.roots.data(), enc_want.data());
        REQUIRE(enc_got == enc_want, "EncodedMatrix differs from the oracle");
        u64 adds, subs, muls; orc_fhe_ntt_counts(&adds, &subs, &muls);
        REQUIRE((u64)server.MulCounter() == muls, "MulCounter %d vs %llu", server.MulCounter(), (unsigned long long)muls);
        std::vector<u64> row(cols), encRow(n), dec(N);
        std::vector<std::vector<u64>> encPlain(rows, std::vector<u64>(n));
        for (int i = 0; i < rows; i++) {
            for (int j = 0; j < cols; j++) row[j] = M[(size_t)i * cols + j];
            orc_core_encode(row.data(), cols, rhoInv, Modulus, field.roots.data(), encPlain[i].data());
        }
        for (int j = 0; j < n; j++) {
            orc_decrypt_decode(P, sk.data(), enc_got.data() + ctw * j, k - 1, 1, dec.data());
            for (int i = 0; i < rows; i++) REQUIRE(dec[i] == encPlain[i][j], "Matrices differ at [%d][%d]", i, j);
        }
        // TestLigeroE2E: MatR / MatZ slot 0 == LigeroProveReference inner products; ciphertexts bit-exact vs oracle
        const size_t l1w = 2 * 2 * (size_t)N;
        std::vector<u64> matR = proof.MatR->Download(), matZ = proof.MatZ->Download(), want(l1w);
        for (int j = 0; j < cols; j++) {
            u64 sc = orc_inner_sum_eval_column(P, cts.data() + ctw * j, rpt.data(), rows, galEls, evkp.data(), ngal, want.data());
            REQUIRE(sc == proof.MatR->Scale, "scale metadata");
            REQUIRE(std::memcmp(want.data(), matR.data() + l1w * j, l1w * 8) == 0, "MatR[%d] differs from the oracle", j);
            orc_inner_sum_eval_column(P, cts.data() + ctw * j, bpt.data(), rows, galEls, evkp.data(), ngal, want.data());
            REQUIRE(std::memcmp(want.data(), matZ.data() + l1w * j, l1w * 8) == 0, "MatZ[%d] differs from the oracle", j);
            unsigned __int128 accR = 0, accZ = 0;
            for (int i = 0; i < rows; i++) {
                accR = (accR + (unsigned __int128)(r[i] % Modulus) * M[(size_t)i * cols + j]) % Modulus;
                accZ = (accZ + (unsigned __int128)b[i] * M[(size_t)i * cols + j]) % Modulus;
            }
            orc_decrypt_decode(P, sk.data(), matR.data() + l1w * j, 1, sc, dec.data());
            REQUIRE(dec[0] == (u64)accR, "MatR[%d] decrypts to %llu, want %llu", j, (unsigned long long)dec[0], (unsigned long long)(u64)accR);
            orc_decrypt_decode(P, sk.data(), matZ.data() + l1w * j, 1, sc, dec.data());
            REQUIRE(dec[0] == (u64)accZ, "MatZ[%d] decrypts wrong", j);
        }
        // queried columns: bit-exact, and Verify's well-formedness relation <col, r> == Encode(MatR)[idx] (fhe/ligero.go:554-563)
        std::vector<u64> qc = proof.QueriedCols->Download(), matRplain(cols), encR(n);
        for (int j = 0; j < cols; j++) {
            unsigned __int128 acc = 0;
            for (int i = 0; i < rows; i++) acc = (acc + (unsigned __int128)(r[i] % Modulus) * M[(size_t)i * cols + j]) % Modulus;
            matRplain[j] = (u64)acc;
        }
        orc_core_encode(matRplain.data(), cols, rhoInv, Modulus, field.roots.data(), encR.data());
        for (int q = 0; q < 24; q++) {
            u64 sc = orc_rescale_to_level1(P, enc_want.data() + ctw * idx[q], want.data());
            REQUIRE(std::memcmp(want.data(), qc.data() + l1w * q, l1w * 8) == 0, "QueriedCols[%d] differs from the oracle", q);
            orc_decrypt_decode(P, sk.data(), qc.data() + l1w * q, 1, sc, dec.data());
            unsigned __int128 acc = 0;
            for (int i = 0; i < rows; i++) acc = (acc + (unsigned __int128)dec[i] * (r[i] % Modulus)) % Modulus;
            REQUIRE((u64)acc == encR[idx[q]], "well-formedness R check failed for column %u", idx[q]);
        }
        // Commit's Merkle span: root and paths against the oracle's core.NewTree over the oracle's leaf buffers
        {
            std::vector<u64> leaf(l1w);
            std::vector<uint8_t> hashes((size_t)n * 32), buf;
            for (int j = 0; j < n; j++) {
                orc_rescale_to_level1(P, enc_want.data() + ctw * j, leaf.data());
                const u64 len = orc_leaf_bytes(leaf.data(), 2, N, lf.header.data(), (uint32_t)lf.header.size(), lf.polyHeader.data(), (uint32_t)lf.polyHeader.size(),
                                               lf.limbHeader.data(), (uint32_t)lf.limbHeader.size(), nullptr);
                buf.resize(len);
                orc_leaf_bytes(leaf.data(), 2, N, lf.header.data(), (uint32_t)lf.header.size(), lf.polyHeader.data(), (uint32_t)lf.polyHeader.size(),
                               lf.limbHeader.data(), (uint32_t)lf.limbHeader.size(), buf.data());
                orc_sha256(buf.data(), len, hashes.data() + 32 * (size_t)j);
            }
            const uint32_t total = orc_merkle_tree(hashes.data(), (uint32_t)n, nullptr);
            std::vector<uint8_t> nodes((size_t)total * 32), path(32 * 40);
            orc_merkle_tree(hashes.data(), (uint32_t)n, nodes.data());
            REQUIRE(proof.Root.size() == 32 && std::memcmp(proof.Root.data(), nodes.data() + 32 * (size_t)(total - 1), 32) == 0, "Merkle root differs from the oracle");
            REQUIRE((int)proof.MerklePaths.size() == ligero.Queries, "one Merkle path per query");
            for (int q = 0; q < 24; q++) {
                const uint32_t d = orc_merkle_path(nodes.data(), (uint32_t)n, idx[q], path.data());
                REQUIRE(proof.MerklePaths[q].size() == d, "path length");
                for (uint32_t l = 0; l < d; l++) REQUIRE(std::memcmp(proof.MerklePaths[q][l].data(), path.data() + 32 * l, 32) == 0, "Merkle path %d level %u", q, l);
            }
            // wire format: EncryptedProof.MarshalBinary against the oracle's WriteTo (fhe/ligero.go:660-704)
            std::vector<uint8_t> wire = proof.MarshalBinary(server, lf), flat;
            for (const auto &p : proof.MerklePaths) for (const auto &hsh : p) flat.insert(flat.end(), hsh.begin(), hsh.end());
            std::vector<uint8_t> wwant(wire.size() + 64);
            const u64 wl = orc_proof_bytes((uint32_t)rows, (uint32_t)cols, (uint32_t)rhoInv, (uint32_t)ligero.Queries, matR.data(), matZ.data(), (uint32_t)cols, 2, N,
                                           lf.header.data(), (uint32_t)lf.header.size(), lf.polyHeader.data(), (uint32_t)lf.polyHeader.size(),
                                           lf.limbHeader.data(), (uint32_t)lf.limbHeader.size(), qc.data(), (uint32_t)ligero.Queries, 2, N,
                                           lf.header.data(), (uint32_t)lf.header.size(), lf.polyHeader.data(), (uint32_t)lf.polyHeader.size(),
                                           lf.limbHeader.data(), (uint32_t)lf.limbHeader.size(), flat.data(), flat.size(), proof.Root.data(), 32, wwant.data());
            REQUIRE(wl == wire.size() && std::memcmp(wire.data(), wwant.data(), wl) == 0, "marshalled proof differs from the oracle (%zu vs %llu bytes)", wire.size(), (unsigned long long)wl);
            bool oob = false;
            try { comm.Tree->GetMerklePath((uint32_t)n); } catch (const Error &) { oob = true; }
            REQUIRE(oob, "GetMerklePath out of bounds must fail");
        }
        // the stages either side: device encoder bit-exact vs the oracle's, device-encrypted columns decrypt (oracle
        // client) to the witness, device decrypt + decode == the oracle's on the prover's level-1 outputs
        {
            auto mform = [](u64 x, u64 q) { return (u64)(((unsigned __int128)x << 64) % q); };
            std::vector<u64> pkm(ctw), skm((size_t)k * N), sn(N);
            for (int c = 0; c < 2; c++) for (int i = 0; i < k; i++) for (int j = 0; j < N; j++)
                pkm[((size_t)c * k + i) * N + j] = mform(pk[((size_t)c * k + i) * N + j], params.Q[i]);
            for (int i = 0; i < k; i++) {
                for (int j = 0; j < N; j++) sn[j] = sk[j] >= 0 ? (u64)sk[j] : params.Q[i] - (u64)(-sk[j]);
                orc_ntt_limb(P, i, sn.data());
                for (int j = 0; j < N; j++) skm[(size_t)i * N + j] = mform(sn[j], params.Q[i]);
            }
            fhe::ClientBFV client(field, params, skm.data(), 0);
            client.LoadPublicKey(pkm.data());
            REQUIRE(client.Encode(r) == rpt && client.Encode(b) == bpt, "device bgv.Encoder.Encode differs from the oracle");
            uint8_t seed32[32]; for (int i = 0; i < 32; i++) seed32[i] = (uint8_t)i;
            fhe::CiphertextStore encd(client, cols);
            client.EncryptColumns((uint32_t)rows, (uint32_t)cols, 0, (uint32_t)cols, seed32, encd);
            std::vector<u64> ec = encd.Download();
            for (int j = 0; j < cols; j++) {
                orc_decrypt_decode(P, sk.data(), ec.data() + ctw * j, k - 1, 1, dec.data());
                for (int i = 0; i < rows; i++) REQUIRE(dec[i] == M[(size_t)i * cols + j], "device-encrypted column %d does not decrypt to the witness", j);
            }
            std::vector<u64> all = client.DecryptDecode(encd, (uint32_t)rows);
            for (int j = 0; j < cols; j++) for (int i = 0; i < rows; i++) REQUIRE(all[(size_t)j * rows + i] == M[(size_t)i * cols + j], "device decrypt of column %d", j);
            fhe::CiphertextStore l1(client, cols, 1);
            l1.Upload(matR.data(), 0, cols); l1.Scale = proof.MatR->Scale;
            std::vector<u64> s0 = client.DecryptDecode(l1, 1);
            for (int j = 0; j < cols; j++) {
                orc_decrypt_decode(P, sk.data(), matR.data() + l1w * j, 1, proof.MatR->Scale, dec.data());
                REQUIRE(s0[j] == dec[0], "device decrypt of MatR[%d] differs from the oracle", j);
            }
        }
        // error behaviour: missing key / bad shapes raise, they do not abort
        fhe::ServerBFV bare(field, params, 0);
        fhe::CiphertextStore one(bare, 1);
        one.Upload(cts.data(), 0, 1);
        bool threw = false;
        try { bare.InnerSum({&one, 0}, 1, rows); } catch (const Error &e) { threw = e.code == LMN_ERR_NOKEY; }
        REQUIRE(threw, "InnerSum without keys must fail with LMN_ERR_NOKEY");
    } catch (const std::exception &e) {
        std::fprintf(stderr, "FAIL: exception: %s\n", e.what());
        return 1;
    }
    orc_params_free(P);
    std::printf("ok: C++ mirror Commit/Prove bit-exact vs oracle and plaintext-correct (logN=%d %dx%d k=%d)\n", logN, rows, cols, k);
    return 0;
}
