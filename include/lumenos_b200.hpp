// lumenos_b200.hpp -- C++ host-side mirror of the reference's `fhe` package surface for the accelerated path, written
// over the C ABI (lumenos_b200.h).  The reference is Go (compiled code) and no Go toolchain exists in this image, so this
// header is the compiled-language host layer: same names, argument meaning and error behaviour as the Go interface
// (ChainSafe/lumenos @ 9f33d62), with Go's `error` returns rendered as exceptions (lumenos::Error):
//
//   fhe::GenerateBGVParamsForNTT   fhe/bfv.go:121-188        fhe::ServerBFV            fhe/bfv.go:13-58
//   fhe::Encode                    fhe/code.go:8-34          fhe::NTT                  fhe/ntt.go:12-17
//   fhe::LigeroCommitter::Commit   fhe/ligero.go:95-124      fhe::LigeroProver::Prove  fhe/ligero.go:194-291
//   fhe::matrixInnerSumEval        fhe/ligero.go:299-370     fhe::calculateQueries     fhe/ligero.go:65-71
//
//   core::MerkleTree / NewTree     core/tree.go:15-21, 76-219  (Commit's "Merkle tree built" span, fhe/ligero.go:105-116)
//   fhe::ServerBFV::EncryptColumns cmd/server/main.go:187-207 fhe::ClientBFV            fhe/bfv.go:60-106, fhe/ligero.go:381-502
//
// Keys, encryption and the merlin transcript stay with the host application (they are Lattigo / Go territory in the
// reference); they enter here as data (plaintext limbs, Galois keys, the zero column, query indices, the framing bytes of
// lattigo's Ciphertext.WriteTo).
#pragma once
#include <cmath>
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "lumenos_b200.h"

namespace lumenos {

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string &m) : std::runtime_error("lumenos_b200: " + m), code(c) {}
};
inline void check(int rc) { if (rc != 0) throw Error(rc, lmn_last_error()); }

namespace fhe {

// ---- bgv.ParametersLiteral as GenerateBGVParamsForNTT returns it ----
struct ParametersLiteral {
    int LogN = 0;
    std::vector<int> LogQ, LogP;
    uint64_t PlaintextModulus = 0;
};
inline ParametersLiteral GenerateBGVParamsForNTT(int nttSize, int logN, uint64_t plaintextModulus) {
    if (nttSize < 2) throw std::invalid_argument("nttSize must be >= 2");
    if (logN <= 0) throw std::invalid_argument("logN must be positive");
    const uint64_t twoN = 2ull << logN;
    if (plaintextModulus % twoN != 1)
        throw std::invalid_argument("plaintextModulus T does not satisfy T = 1 (mod 2N)");
    int bits = 64 - __builtin_clzll(plaintextModulus);
    int k = __builtin_ctz((unsigned)nttSize) + (bits > 45 ? 0 : -2);
    ParametersLiteral lit;
    lit.LogN = logN; lit.PlaintextModulus = plaintextModulus;
    for (int i = 0; i < k; i++) lit.LogQ.push_back(i == 0 ? 58 : 56);
    lit.LogP = {55, 55};
    return lit;
}
inline int calculateQueries(double securityBits, int rhoInv) {
    double term = std::log2(1.0 + 1.0 / rhoInv);
    if (1.0 - term <= 0) return 0;
    return (int)std::ceil(securityBits / (1.0 - term));
}

// ---- bgv.Parameters as the host holds them ----
struct Parameters {
    int LogN = 0;
    std::vector<uint64_t> Q, P, PsiQ, PsiP;   // moduli and the primitive 2N-th roots the host rings use (plain form)
    uint64_t T = 0;
    uint64_t PsiT = 0;                        // primitive 2N-th root of the encoder's ring T (Encode / Encrypt / Decode only)
    int N() const { return 1 << LogN; }
    int MaxLevel() const { return (int)Q.size() - 1; }
    uint64_t GaloisElement(int k) const {
        uint64_t mask = 2ull * N() - 1, g = 1, b = 5;
        for (unsigned e = (unsigned)k & (unsigned)(N() / 2 - 1); e; e >>= 1) { if (e & 1) g = (g * b) & mask; b = (b * b) & mask; }
        return g;
    }
    std::vector<uint64_t> GaloisElementsForInnerSum(int batch, int n) const {
        // lattigo: rotations batch * {2^i < n} and batch * n (identity elements included: it generates keys for them and the library
        // ignores them), plus the row swap when n * batch > N / 2
        std::vector<uint64_t> g;
        for (int s = 1; s < n; s <<= 1) g.push_back(GaloisElement(s * batch));
        g.push_back(GaloisElement(n * batch));
        if (n * batch > N() / 2) g.push_back(2ull * N() - 1);
        return g;
    }
};

// core.PrimeField as far as the path reads it (core/field.go:29-47)
struct PrimeField {
    uint64_t modulus = 0;
    std::vector<uint64_t> roots;   // RootsForward, Montgomery form, N() entries
    int N() const { return (int)roots.size(); }
    uint64_t Modulus() const { return modulus; }
    uint64_t RootForwardUint64(int i) const { return roots.at(i); }
};

class ServerBFV;

// device-resident []*rlwe.Ciphertext
class CiphertextStore {
  public:
    CiphertextStore(ServerBFV &be, uint32_t count, int level = -1);
    ~CiphertextStore() { if (h_) lmn_store_destroy(h_); }
    CiphertextStore(const CiphertextStore &) = delete;
    CiphertextStore &operator=(const CiphertextStore &) = delete;
    lmn_store *handle() const { return h_; }
    uint32_t size() const { return count_; }
    int Level() const { return level_; }
    size_t words() const { return 2 * (size_t)(level_ + 1) * n_; }
    void Upload(const uint64_t *cts, uint32_t first, uint32_t count) { check(lmn_store_upload(h_, first, count, cts)); }
    void Download(uint64_t *cts, uint32_t first, uint32_t count) const { check(lmn_store_download(h_, first, count, cts)); }
    std::vector<uint64_t> Download() const { std::vector<uint64_t> v(words() * count_); Download(v.data(), 0, count_); return v; }
    uint64_t Scale = 1;   // rlwe.MetaData.Scale (mod T), tracked by the host as in the reference

  private:
    lmn_store *h_ = nullptr;
    uint32_t count_ = 0;
    int level_ = 0;
    size_t n_ = 0;
};
using Operand = std::pair<CiphertextStore *, uint32_t>;   // one ciphertext: (store, index)

class ServerBFV {
  public:
    ServerBFV(const PrimeField &ptField, const Parameters &params, int device = 0) : ptField_(ptField), params_(params) {
        lmn_params p{};
        p.log_n = (uint32_t)params.LogN; p.num_q = (uint32_t)params.Q.size(); p.num_p = (uint32_t)params.P.size();
        p.q = params.Q.data(); p.p = params.P.data(); p.psi_q = params.PsiQ.data(); p.psi_p = params.PsiP.data(); p.t = params.T;
        p.scalar_centered = p.moddown_round_nearest = p.innersum_rowswap_last = 0xffffffffu;
        check(lmn_ctx_create(&p, device, &ctx_));
    }
    ~ServerBFV() {
        for (auto &pl : plans_) lmn_plan_destroy(pl.second);
        if (ctx_) lmn_ctx_destroy(ctx_);
    }
    ServerBFV(const ServerBFV &) = delete;
    ServerBFV &operator=(const ServerBFV &) = delete;

    lmn_ctx *ctx() const { return ctx_; }
    const Parameters &params() const { return params_; }
    const PrimeField &Field() const { return ptField_; }
    int MulCounter() const { return mulCounter_; }
    void LoadGaloisKey(uint64_t galEl, const uint64_t *evk) { check(lmn_load_galois_key(ctx_, galEl, evk)); }
    void Synchronize() { check(lmn_ctx_synchronize(ctx_)); }

    // promoted bgv.Evaluator methods used by the path
    void Add(Operand a, Operand b, Operand o) { check(lmn_ct_add(ctx_, a.first->handle(), a.second, b.first->handle(), b.second, o.first->handle(), o.second)); }
    void Sub(Operand a, Operand b, Operand o) { check(lmn_ct_sub(ctx_, a.first->handle(), a.second, b.first->handle(), b.second, o.first->handle(), o.second)); }
    void Mul(Operand a, uint64_t scalar, Operand o) { mulCounter_++; check(lmn_ct_mul_scalar(ctx_, a.first->handle(), a.second, scalar, o.first->handle(), o.second)); }
    void MulNew(Operand a, const uint64_t *plaintext, Operand o) { mulCounter_++; check(lmn_ct_mul_pt(ctx_, a.first->handle(), a.second, plaintext, o.first->handle(), o.second)); }
    void InnerSum(Operand a, int batchSize, int n) {
        if (batchSize != 1) throw std::invalid_argument("the path uses InnerSum(ct, 1, rows) only (fhe/ligero.go:325)");
        check(lmn_ct_inner_sum(ctx_, a.first->handle(), a.second, (uint32_t)n));
    }
    void Rescale(Operand a, Operand o) { check(lmn_ct_rescale(ctx_, a.first->handle(), a.second, o.first->handle(), o.second)); }
    void CopyNew(Operand a, Operand o) { check(lmn_ct_copy(ctx_, a.first->handle(), a.second, o.first->handle(), o.second)); }
    uint64_t RescaleScale(int from, int to) const { return lmn_rescale_scale(ctx_, (uint32_t)from, (uint32_t)to); }

    // the bgv.Encoder / rlwe.Encryptor halves of ServerBFV (fhe/bfv.go:13-27) on the device
    void LoadPlaintextRing() { check(lmn_load_plaintext_ring(ctx_, params_.PsiT)); }
    void LoadPublicKey(const uint64_t *pk) { LoadPlaintextRing(); check(lmn_load_public_key(ctx_, pk)); }   // [2][k][N] NTT + Montgomery
    std::vector<uint64_t> Encode(const std::vector<uint64_t> &values) {                                      // bgv.Encoder.Encode at max level
        std::vector<uint64_t> pt(params_.Q.size() * (size_t)params_.N());
        check(lmn_encode_plaintext(ctx_, values.data(), (uint32_t)values.size(), pt.data()));
        return pt;
    }
    // "Encrypt matrix" (cmd/server/main.go:187-207); values == nullptr: core.RandomMatrixRowMajor's ChaCha20 witness on the device
    void EncryptColumns(uint32_t rows, uint32_t cols, uint32_t firstCol, uint32_t count, const uint8_t encSeed[32], CiphertextStore &out,
                        const uint64_t *valuesColMajor = nullptr);
    void EncryptZero(const uint8_t encSeed[32], CiphertextStore &out, uint32_t index = 0, uint32_t streamId = 0xffffffffu);

    lmn_plan *plan(uint32_t cols, uint32_t rhoInv) {
        auto key = std::make_pair(cols, rhoInv);
        for (auto &pl : plans_) if (pl.first == key) return pl.second;
        if ((uint32_t)ptField_.N() != cols * rhoInv) throw std::invalid_argument("plaintext field degree does not match the codeword length");
        lmn_plan *p = nullptr;
        check(lmn_plan_create(ctx_, cols, rhoInv, ptField_.roots.data(), (uint32_t)ptField_.roots.size(), &p));
        plans_.emplace_back(key, p);
        return p;
    }
    void addMuls(uint64_t m) { mulCounter_ += (int)m; }

  private:
    PrimeField ptField_;
    Parameters params_;
    lmn_ctx *ctx_ = nullptr;
    int mulCounter_ = 0;
    std::vector<std::pair<std::pair<uint32_t, uint32_t>, lmn_plan *>> plans_;
};

inline void ServerBFV::EncryptColumns(uint32_t rows, uint32_t cols, uint32_t firstCol, uint32_t count, const uint8_t encSeed[32], CiphertextStore &out,
                                      const uint64_t *valuesColMajor) {
    uint8_t key[32] = {1};                   // core/utils.go:55-56: seed = LE64(1) followed by zeros
    check(lmn_encrypt_columns(ctx_, valuesColMajor, valuesColMajor ? nullptr : key, rows, cols, firstCol, count, encSeed, out.handle()));
}
inline void ServerBFV::EncryptZero(const uint8_t encSeed[32], CiphertextStore &out, uint32_t index, uint32_t streamId) {
    check(lmn_encrypt_zero(ctx_, encSeed, streamId, out.handle(), index));
}

// fhe.ClientBFV as far as EncryptedProof.Decrypt uses it (fhe/bfv.go:60-106, fhe/ligero.go:381-502, 577-636): Decrypt + Decode on the device
class ClientBFV : public ServerBFV {
  public:
    ClientBFV(const PrimeField &ptField, const Parameters &params, const uint64_t *sk, int device = 0) : ServerBFV(ptField, params, device) {
        LoadPlaintextRing();
        check(lmn_load_secret_key(ctx(), sk));      // [k][N] = sk.Value.Q, NTT + Montgomery
    }
    // the first nValues slots of store[first .. first + count): [count][nValues]
    std::vector<uint64_t> DecryptDecode(const CiphertextStore &s, uint32_t nValues, uint32_t first = 0, uint32_t count = 0) {
        if (!count) count = s.size() - first;
        std::vector<uint64_t> out((size_t)count * nValues);
        check(lmn_decrypt_decode(ctx(), s.handle(), first, count, s.Scale, nValues, out.data()));
        return out;
    }
};

inline CiphertextStore::CiphertextStore(ServerBFV &be, uint32_t count, int level)
    : count_(count), level_(level < 0 ? be.params().MaxLevel() : level), n_((size_t)be.params().N()) {
    check(lmn_store_create(be.ctx(), count, (uint32_t)level_, &h_));
}

// fhe.NTT (fhe/ntt.go:12-17), len(values) == size
inline std::unique_ptr<CiphertextStore> NTT(CiphertextStore &values, int size, ServerBFV &backend) {
    if ((int)values.size() != size) throw std::invalid_argument("this build evaluates whole-slice transforms: len(values) must equal size");
    auto out = std::make_unique<CiphertextStore>(backend, (uint32_t)size);
    lmn_plan *pl = backend.plan((uint32_t)size, 1);
    check(lmn_ntt(backend.ctx(), pl, values.handle(), out->handle()));
    uint64_t a, s, m; check(lmn_plan_counts(pl, &a, &s, &m)); backend.addMuls(m);
    return out;
}
// fhe.Encode (fhe/code.go:8-34); zeroCol = the host-produced encryption of the zero vector
inline std::unique_ptr<CiphertextStore> Encode(CiphertextStore &matrix, int rows, int rhoInv, ServerBFV &backend, Operand zeroCol) {
    (void)rows;
    auto encoded = std::make_unique<CiphertextStore>(backend, matrix.size() * (uint32_t)rhoInv);
    lmn_plan *pl = backend.plan(matrix.size(), (uint32_t)rhoInv);
    check(lmn_encode(backend.ctx(), pl, matrix.handle(), zeroCol.first->handle(), zeroCol.second, encoded->handle()));
    uint64_t a, s, m; check(lmn_plan_counts(pl, &a, &s, &m)); backend.addMuls(m);
    return encoded;
}
// matrixInnerSumEval (fhe/ligero.go:299-370); plaintext = [k][N] NTT-domain limbs of the bgv plaintext
inline std::unique_ptr<CiphertextStore> matrixInnerSumEval(CiphertextStore &matrix, const uint64_t *plaintext, int rows, ServerBFV &backend) {
    const int lout = std::min(1, backend.params().MaxLevel());
    auto out = std::make_unique<CiphertextStore>(backend, matrix.size(), lout);
    check(lmn_inner_product(backend.ctx(), matrix.handle(), plaintext, (uint32_t)rows, out->handle()));
    out->Scale = backend.RescaleScale(backend.params().MaxLevel(), lout);
    return out;
}

struct LigeroMetadata { int Rows = 0, Cols = 0, RhoInv = 0, Queries = 0; };

typedef std::vector<std::vector<uint8_t>> MerklePath;   // core/tree.go:33

// framing of the leaf byte stream ct.WriteTo(buf) (fhe/ligero.go:157-158); see lmn_leaf_format
struct LeafFormat {
    std::vector<uint8_t> header, polyHeader, limbHeader;
    lmn_leaf_format c() const {
        return lmn_leaf_format{header.data(), (uint32_t)header.size(), polyHeader.data(), (uint32_t)polyHeader.size(), limbHeader.data(), (uint32_t)limbHeader.size()};
    }
};

struct EncryptedProof {   // fhe/ligero.go:185-192
    LigeroMetadata Metadata;
    std::unique_ptr<CiphertextStore> MatR, MatZ, QueriedCols;
    std::vector<MerklePath> MerklePaths;
    std::vector<uint8_t> Root;

    // MarshalBinary / WriteTo (fhe/ligero.go:646-704): the ciphertext streams are serialised on the GPU straight into the result.
    // matFormat frames MatR / MatZ, queryFormat (default: the same) the queried columns (they differ after a ring switch).
    std::vector<uint8_t> MarshalBinary(ServerBFV &backend, const LeafFormat &matFormat, const LeafFormat *queryFormat = nullptr) const;
};
// core.MerkleTree (core/tree.go:15-21): built on the GPU by core.NewTree's rule, node hashes on the host
class MerkleTree {
  public:
    MerkleTree(ServerBFV &be, CiphertextStore &leaves, const LeafFormat &f) { lmn_leaf_format c = f.c(); check(lmn_merkle_build(be.ctx(), leaves.handle(), &c, &h_)); }
    ~MerkleTree() { if (h_) lmn_merkle_destroy(h_); }
    MerkleTree(const MerkleTree &) = delete;
    MerkleTree &operator=(const MerkleTree &) = delete;
    std::vector<uint8_t> MerkleRoot() const { std::vector<uint8_t> r(32); check(lmn_merkle_root(h_, r.data())); return r; }
    MerklePath GetMerklePath(uint32_t index) const {
        const uint32_t d = lmn_merkle_depth(h_);
        std::vector<uint8_t> flat(32 * (size_t)(d ? d : 1));
        check(lmn_merkle_path(h_, index, flat.data()));     // "index %d out of bounds" -> Error, as tree.go:181-183
        MerklePath p(d);
        for (uint32_t i = 0; i < d; i++) p[i].assign(flat.begin() + 32 * i, flat.begin() + 32 * (i + 1));
        return p;
    }
  private:
    lmn_merkle *h_ = nullptr;
};

inline std::vector<uint8_t> EncryptedProof::MarshalBinary(ServerBFV &backend, const LeafFormat &matFormat, const LeafFormat *queryFormat) const {
    const lmn_ligero_metadata md{(uint32_t)Metadata.Rows, (uint32_t)Metadata.Cols, (uint32_t)Metadata.RhoInv, (uint32_t)Metadata.Queries};
    const lmn_leaf_format fm = matFormat.c(), fq = queryFormat ? queryFormat->c() : fm;
    std::vector<uint8_t> paths;
    for (const MerklePath &p : MerklePaths) for (const auto &hsh : p) paths.insert(paths.end(), hsh.begin(), hsh.end());   // MerklePath.WriteTo, core/tree.go:270-277
    size_t n = 0;
    check(lmn_proof_marshal(backend.ctx(), &md, MatR->handle(), MatZ->handle(), QueriedCols->handle(), &fm, &fq, paths.data(), paths.size(),
                            Root.data(), Root.size(), nullptr, 0, &n));
    std::vector<uint8_t> out(n);
    check(lmn_proof_marshal(backend.ctx(), &md, MatR->handle(), MatZ->handle(), QueriedCols->handle(), &fm, &fq, paths.data(), paths.size(),
                            Root.data(), Root.size(), out.data(), out.size(), &n));
    return out;
}

class LigeroCommitter;
class LigeroProver {   // fhe/ligero.go:32-37
  public:
    LigeroMetadata Committer;
    CiphertextStore *Matrix = nullptr;
    std::unique_ptr<CiphertextStore> EncodedMatrix;
    std::unique_ptr<CiphertextStore> Leaves;     // level-1 leaves of Commit, kept on the device (query stage = gather)
    std::unique_ptr<MerkleTree> Tree;
    std::vector<uint8_t> Root;

    // Prove (fhe/ligero.go:194-291): rPt, bPt = limbs of backend.Encode(r), backend.Encode(b); queryIndices from the transcript
    EncryptedProof Prove(const uint64_t *rPt, const uint64_t *bPt, const std::vector<uint32_t> &queryIndices, ServerBFV &backend) {
        if ((int)queryIndices.size() != Committer.Queries) throw std::invalid_argument("unexpected number of query indices");
        const int lout = std::min(1, backend.params().MaxLevel());
        EncryptedProof proof;
        proof.Metadata = Committer;
        proof.MatR = std::make_unique<CiphertextStore>(backend, Matrix->size(), lout);
        proof.MatZ = std::make_unique<CiphertextStore>(backend, Matrix->size(), lout);
        check(lmn_inner_product2(backend.ctx(), Matrix->handle(), rPt, bPt, (uint32_t)Committer.Rows, proof.MatR->handle(), proof.MatZ->handle()));
        proof.QueriedCols = std::make_unique<CiphertextStore>(backend, (uint32_t)queryIndices.size(), lout);
        if (Leaves) check(lmn_store_gather(backend.ctx(), Leaves->handle(), queryIndices.data(), (uint32_t)queryIndices.size(), proof.QueriedCols->handle()));
        else check(lmn_query_columns(backend.ctx(), EncodedMatrix->handle(), queryIndices.data(), (uint32_t)queryIndices.size(), proof.QueriedCols->handle()));
        if (Tree) { for (uint32_t ix : queryIndices) proof.MerklePaths.push_back(Tree->GetMerklePath(ix)); proof.Root = Root; }   // ligero.go:275-279
        proof.MatR->Scale = proof.MatZ->Scale = proof.QueriedCols->Scale = backend.RescaleScale(backend.params().MaxLevel(), lout);
        return proof;
    }
};

class LigeroCommitter : public LigeroMetadata {   // fhe/ligero.go:26-63
  public:
    LigeroCommitter(double securityBits, int rows, int cols, int rhoInv) {
        if ((long long)rows * cols <= 0) throw std::invalid_argument("size must be positive");
        if (securityBits <= 0) throw std::invalid_argument("securityBits must be positive");
        Rows = rows; Cols = cols; RhoInv = rhoInv; Queries = calculateQueries(securityBits, rhoInv);
    }
    // Commit (fhe/ligero.go:95-124): the Encode span and, when a leaf framing is given, the "Merkle tree built" span
    // (processLeafParallel :126-183 + core.NewTree); the reference's second return value is prover.Root
    LigeroProver Commit(CiphertextStore &matrix, ServerBFV &backend, Operand zeroCol, const LeafFormat *leafFormat = nullptr) {
        if ((int)matrix.size() != Cols) throw std::invalid_argument("matrix must hold one ciphertext per column");
        LigeroProver pr;
        pr.Committer = *this;
        pr.Matrix = &matrix;
        pr.EncodedMatrix = Encode(matrix, Rows, RhoInv, backend, zeroCol);
        if (leafFormat) {
            const int lout = std::min(1, backend.params().MaxLevel());
            pr.Leaves = std::make_unique<CiphertextStore>(backend, pr.EncodedMatrix->size(), lout);
            check(lmn_commit_leaves(backend.ctx(), pr.EncodedMatrix->handle(), pr.Leaves->handle()));
            pr.Leaves->Scale = backend.RescaleScale(backend.params().MaxLevel(), lout);
            pr.Tree = std::make_unique<MerkleTree>(backend, *pr.Leaves, *leafFormat);
            pr.Root = pr.Tree->MerkleRoot();
        }
        return pr;
    }
};

}  // namespace fhe
}  // namespace lumenos
