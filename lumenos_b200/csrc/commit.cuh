// Leaf byte-stream description and the launchers of kernels_commit.cu (Merkle commitment over the level-1 leaves).
#pragma once
#include "kernels.cuh"

namespace lmn {

constexpr int LEAF_HDR_MAX = 16;   // bytes of a poly / limb header held inline

struct LeafFormat {
    const uint8_t *hdr;            // device copy of the leaf header (lattigo: MetaData bytes + the vector length)
    uint32_t hdr_len;
    uint8_t poly_hdr[LEAF_HDR_MAX]; uint32_t poly_hdr_len;     // before each of the 2 polynomials
    uint8_t limb_hdr[LEAF_HDR_MAX]; uint32_t limb_hdr_len;     // before each limb
    uint32_t nl, logn;             // limbs per polynomial (level + 1), log2 N
    uint32_t limb_len, poly_len;   // limb_hdr_len + 8 N ; poly_hdr_len + nl * limb_len
    uint32_t total_len;            // hdr_len + 2 poly_len
    uint32_t padded_len;           // SHA-256 padded: multiple of 64
};

void launch_leaf_serialize(const Launch &L, const LeafFormat &F, const u64 *cts, uint64_t ct_words, uint32_t n_leaves, u64 *out);
// bytes [byte0, byte_end) of the unpadded concatenation of the ciphertext streams of `cts`, packed into u64 words
void launch_ct_stream(const Launch &L, const LeafFormat &F, const u64 *cts, uint64_t ct_words, uint64_t byte0, uint64_t byte_end, u64 *out);
void launch_sha256_leaves(const Launch &L, const u64 *ser, uint32_t padded_len, uint32_t n_leaves, uint32_t *hashes);
void launch_sha256_nodes(const Launch &L, const uint32_t *children, uint32_t n_children, uint32_t *parents);

}  // namespace lmn
