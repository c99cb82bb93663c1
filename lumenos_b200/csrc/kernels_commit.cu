// Commit eval after Encode: the Merkle tree over the level-1 leaves.
//   reference: fhe/ligero.go:105-116 (Commit), :126-183 (processLeafParallel: CopyNew, Rescale to level 1, ct.WriteTo),
//              core/tree.go:37-74 (calculateNodeHash: SHA-256 of the leaf bytes / of left || right), :76-163 (NewTree, an odd
//              node is paired with itself), :176-219 (GetMerklePath).
// The rescale of the n encoded ciphertexts is the rescale pipeline of the prove path (kernels_blk.cu); here:
//   k_leaf_serialize   leaf byte stream = header || 2 x ( poly_header || (level+1) x ( limb_header || N little-endian u64 ) )
//                      plus the SHA-256 padding, written as whole 64-byte blocks (HBM-bound, byte-exact for any header length)
//   k_sha256_leaves    one lane per leaf (a leaf is one sequential SHA-256 chain; n chains), a producer warp expanding the message
//                      schedule for the consumer warp that runs the rounds
//   k_sha256_nodes     one thread per parent: SHA-256(left || right), right = left for the unpaired last node
// The framing bytes are the caller's (lattigo's rlwe.Ciphertext.WriteTo layout is owned by the Go host).
#include <cstdlib>
#include "kernels.cuh"
#include "commit.cuh"

namespace lmn {

__constant__ uint32_t SHA_K[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5, 0xd807aa98, 0x12835b01, 0x243185be,
    0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174, 0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa,
    0x5cb0a9dc, 0x76f988da, 0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967, 0x27b70a85,
    0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85, 0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3,
    0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070, 0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f,
    0x682e6ff3, 0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};

__device__ __forceinline__ uint32_t rotr(uint32_t x, int n) { return __funnelshift_r(x, x, n); }
__device__ __forceinline__ uint32_t bswap32(uint32_t x) { return __byte_perm(x, 0, 0x0123); }

// one compression: w[16] = the block as big-endian words (destroyed), st[8] updated
__device__ __forceinline__ void sha256_compress(uint32_t (&st)[8], uint32_t (&w)[16]) {
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
#pragma unroll
    for (int i = 0; i < 64; i++) {
        if (i >= 16) {
            const uint32_t w15 = w[(i + 1) & 15], w2 = w[(i + 14) & 15];
            const uint32_t s0 = rotr(w15, 7) ^ rotr(w15, 18) ^ (w15 >> 3);
            const uint32_t s1 = rotr(w2, 17) ^ rotr(w2, 19) ^ (w2 >> 10);
            w[i & 15] = w[i & 15] + s0 + w[(i + 9) & 15] + s1;
        }
        const uint32_t S1 = rotr(e, 6) ^ rotr(e, 11) ^ rotr(e, 25);
        const uint32_t ch = (e & f) ^ (~e & g);
        const uint32_t t1 = h + S1 + ch + SHA_K[i] + w[i & 15];
        const uint32_t S0 = rotr(a, 2) ^ rotr(a, 13) ^ rotr(a, 22);
        const uint32_t mj = (a & b) ^ (a & c) ^ (b & c);
        const uint32_t t2 = S0 + mj;
        h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}
__device__ __forceinline__ void sha256_init(uint32_t (&st)[8]) {
    st[0] = 0x6a09e667; st[1] = 0xbb67ae85; st[2] = 0x3c6ef372; st[3] = 0xa54ff53a;
    st[4] = 0x510e527f; st[5] = 0x9b05688c; st[6] = 0x1f83d9ab; st[7] = 0x5be0cd19;
}

// ---------------------------------------------------------------------------------------------------
// byte `pos` of the (unpadded) leaf stream of one ciphertext `ct` = [2][nl][N] words
__device__ __forceinline__ uint32_t leaf_byte(const LeafFormat &F, const u64 *__restrict__ ct, uint32_t pos) {
    if (pos < F.hdr_len) return F.hdr[pos];
    pos -= F.hdr_len;
    const uint32_t poly = pos / F.poly_len, r = pos - poly * F.poly_len;
    if (r < F.poly_hdr_len) return F.poly_hdr[r];
    const uint32_t r2 = r - F.poly_hdr_len;
    const uint32_t limb = r2 / F.limb_len, s = r2 - limb * F.limb_len;
    if (s < F.limb_hdr_len) return F.limb_hdr[s];
    const uint32_t b = s - F.limb_hdr_len;
    const u64 w = ct[(((size_t)poly * F.nl + limb) << F.logn) + (b >> 3)];
    return (uint32_t)(w >> (8 * (b & 7))) & 0xffu;
}

// word w of the padded stream of one leaf: bytes 8w..8w+7 (little-endian packing of the byte sequence into the u64)
__device__ __forceinline__ u64 leaf_word(const LeafFormat &F, const u64 *__restrict__ ct, uint32_t w) {
    const uint32_t pos = 8 * w;
    if (pos >= F.hdr_len && pos + 8 <= F.total_len) {
        // fast path: the 8 bytes lie inside the coefficients of one limb
        const uint32_t p0 = pos - F.hdr_len;
        const uint32_t poly = p0 / F.poly_len, r = p0 - poly * F.poly_len;
        if (r >= F.poly_hdr_len) {
            const uint32_t r2 = r - F.poly_hdr_len;
            const uint32_t limb = r2 / F.limb_len, s = r2 - limb * F.limb_len;
            if (s >= F.limb_hdr_len && s + 8 <= F.limb_len) {
                const uint32_t b = s - F.limb_hdr_len, sh = 8 * (b & 7);
                const u64 *src = ct + (((size_t)poly * F.nl + limb) << F.logn) + (b >> 3);
                u64 v = src[0] >> sh;
                if (sh) v |= src[1] << (64 - sh);
                return v;
            }
        }
    }
    u64 v = 0;
#pragma unroll 1
    for (int i = 0; i < 8; i++) {
        const uint32_t p = pos + i;
        uint32_t byte;
        if (p < F.total_len) byte = leaf_byte(F, ct, p);
        else if (p == F.total_len) byte = 0x80;
        else if (p < F.padded_len - 8) byte = 0;
        else byte = (uint32_t)((8ull * F.total_len) >> (8 * (F.padded_len - 1 - p))) & 0xffu;   // big-endian bit length
        v |= (u64)byte << (8 * i);
    }
    return v;
}
// out[leaf][2w], out[leaf][2w+1]: two words (one 16-byte store) per thread; the padded length is a multiple of 64 bytes
// (the leaf index is folded into blockIdx.x - bx blocks per leaf - so that any leaf count lmn_plan_create admits fits the grid)
__global__ void __launch_bounds__(256) k_leaf_serialize(const LeafFormat F, const u64 *__restrict__ cts, uint64_t ct_words,
                                                        u64 *__restrict__ out, uint32_t bx) {
    const uint32_t leaf = blockIdx.x / bx;
    const uint32_t w2 = 2 * ((blockIdx.x - leaf * bx) * blockDim.x + threadIdx.x);
    const uint32_t words = F.padded_len / 8;
    if (w2 >= words) return;
    const u64 *ct = cts + (size_t)leaf * ct_words;
    ulonglong2 v;
    v.x = leaf_word(F, ct, w2);
    v.y = leaf_word(F, ct, w2 + 1);
    reinterpret_cast<ulonglong2 *>(out + (size_t)leaf * words)[w2 >> 1] = v;
}

// Wire format of the proof (EncryptedProof.WriteTo, fhe/ligero.go:660-691): MatR, MatZ and QueriedCols are written as
// ct.WriteTo(cts[0]) || ct.WriteTo(cts[1]) || ... without padding.  out[w] = bytes [byte0 + 8w, byte0 + 8w + 8) of that
// concatenation (little-endian packing), bytes at or beyond byte_end are zero.  A stream need not be a multiple of 8 bytes long, so
// a word may straddle two ciphertexts: those words (and the ones touching framing bytes) take the byte-wise path.
__global__ void __launch_bounds__(256) k_ct_stream(const LeafFormat F, const u64 *__restrict__ cts, uint64_t ct_words, uint64_t byte0,
                                                   uint64_t byte_end, u64 *__restrict__ out) {
    const uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t p = byte0 + 8 * w;
    if (p >= byte_end) return;
    const uint64_t ci = p / F.total_len;
    const uint32_t pos = (uint32_t)(p - ci * F.total_len);
    u64 v = 0;
    bool done = false;
    if (pos >= F.hdr_len && pos + 8 <= F.total_len) {
        const uint32_t p0 = pos - F.hdr_len;
        const uint32_t poly = p0 / F.poly_len, r = p0 - poly * F.poly_len;
        if (r >= F.poly_hdr_len) {
            const uint32_t r2 = r - F.poly_hdr_len;
            const uint32_t limb = r2 / F.limb_len, s = r2 - limb * F.limb_len;
            if (s >= F.limb_hdr_len && s + 8 <= F.limb_len) {
                const uint32_t b = s - F.limb_hdr_len, sh = 8 * (b & 7);
                const u64 *src = cts + ci * ct_words + (((size_t)poly * F.nl + limb) << F.logn) + (b >> 3);
                v = src[0] >> sh;
                if (sh) v |= src[1] << (64 - sh);
                done = true;
            }
        }
    }
    if (!done) {
#pragma unroll 1
        for (int i = 0; i < 8; i++) {
            const uint64_t pp = p + i;
            if (pp >= byte_end) break;
            const uint64_t c2 = pp / F.total_len;
            v |= (u64)leaf_byte(F, cts + c2 * ct_words, (uint32_t)(pp - c2 * F.total_len)) << (8 * i);
        }
    }
    out[w] = v;
}

// SHA-256 of the serialised leaves.  A leaf is ONE message (one sequential chain of compressions), so the only parallelism is across
// leaves: n chains, one per lane.  What can be taken off the chain is the message schedule: a CTA is a pair of warps over the same
// 32 leaves - the PRODUCER warp loads each 64-byte block, expands W[0..63] and leaves W[i] + K[i] in a 4-deep ring in shared memory,
// the CONSUMER warp runs only the 64 rounds (per round: SHF, LOP3, IADD3 on the e-chain).  The two warps sit on different
// schedulers and hand slots over with named barriers (bar.arrive releases the producer's stores, bar.sync acquires them).
// Measured on B200, 8192 leaves of 512 KiB: 15.8 ms with one thread doing both halves -> see DESIGN.md section 8.
constexpr int SHA_DEPTH = 4;
__device__ __forceinline__ void named_sync(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void named_arrive(int id) { asm volatile("bar.arrive %0, 64;" ::"r"(id) : "memory"); }
constexpr int SHA_PRODUCERS = 1;     // producer warps per consumer warp (2, taking alternate blocks: 10.2 ms against 8.3 - the consumer is the limit: 960 instructions per block at one issue per two cycles per warp)
__global__ void __launch_bounds__(32 * (1 + SHA_PRODUCERS)) k_sha256_leaves(const uint4 *__restrict__ ser, uint32_t blocks_per_leaf, uint32_t n_leaves,
                                                      uint32_t *__restrict__ hashes) {
    __shared__ uint32_t ring[SHA_DEPTH][64][32];      // [slot][round][lane] = W + K
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t leaf_id = blockIdx.x * 32 + lane;
    const uint32_t leaf = leaf_id < n_leaves ? leaf_id : n_leaves - 1;     // idle lanes shadow the last leaf (barriers count whole warps)
    if (warp >= 1) {
        const uint4 *p = ser + (size_t)leaf * blocks_per_leaf * 4;
        const uint32_t b0 = (uint32_t)(warp - 1);
        uint4 nx[4];
        if (b0 < blocks_per_leaf) {
#pragma unroll
            for (int j = 0; j < 4; j++) nx[j] = __ldcs(p + 4 * (size_t)b0 + j);
        }
#pragma unroll 1
        for (uint32_t b = b0; b < blocks_per_leaf; b += SHA_PRODUCERS) {
            uint32_t w[16];
#pragma unroll
            for (int j = 0; j < 4; j++) { w[4 * j] = bswap32(nx[j].x); w[4 * j + 1] = bswap32(nx[j].y); w[4 * j + 2] = bswap32(nx[j].z); w[4 * j + 3] = bswap32(nx[j].w); }
            if (b + SHA_PRODUCERS < blocks_per_leaf) {
#pragma unroll
                for (int j = 0; j < 4; j++) nx[j] = __ldcs(p + 4 * (size_t)(b + SHA_PRODUCERS) + j);
            }
            const int slot = (int)(b & (SHA_DEPTH - 1));
            if (b >= SHA_DEPTH) named_sync(1 + SHA_DEPTH + slot);           // the consumer has finished with this slot
#pragma unroll
            for (int i = 0; i < 64; i++) {
                if (i >= 16) {
                    const uint32_t w15 = w[(i + 1) & 15], w2 = w[(i + 14) & 15];
                    const uint32_t s0 = rotr(w15, 7) ^ rotr(w15, 18) ^ (w15 >> 3);
                    const uint32_t s1 = rotr(w2, 17) ^ rotr(w2, 19) ^ (w2 >> 10);
                    w[i & 15] = w[i & 15] + s0 + w[(i + 9) & 15] + s1;
                }
                ring[slot][i][lane] = w[i & 15] + SHA_K[i];
            }
            named_arrive(1 + slot);                                        // slot full
        }
    } else {
        uint32_t st[8];
        sha256_init(st);
#pragma unroll 1
        for (uint32_t b = 0; b < blocks_per_leaf; b++) {
            const int slot = (int)(b & (SHA_DEPTH - 1));
            named_sync(1 + slot);
            uint32_t a = st[0], bb = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
#pragma unroll
            for (int i = 0; i < 64; i++) {
                // the chain through e is SHF -> LOP3 -> IADD3 per round: h + W + K and d are three rounds old, so x and y are off it
                const uint32_t x = h + ring[slot][i][lane];
                const uint32_t y = x + d;
                const uint32_t S1 = rotr(e, 6) ^ rotr(e, 11) ^ rotr(e, 25);
                const uint32_t ch = (e & f) ^ (~e & g);
                const uint32_t en = y + S1 + ch;                          // = d + t1
                const uint32_t t1 = x + S1 + ch;
                const uint32_t S0 = rotr(a, 2) ^ rotr(a, 13) ^ rotr(a, 22);
                const uint32_t mj = (a & bb) ^ (a & c) ^ (bb & c);
                h = g; g = f; f = e; e = en; d = c; c = bb; bb = a; a = t1 + S0 + mj;
            }
            if (b + SHA_DEPTH < blocks_per_leaf) named_arrive(1 + SHA_DEPTH + slot);   // slot free (only while the producer still waits for it)
            st[0] += a; st[1] += bb; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
        }
        if (leaf_id < n_leaves) {
#pragma unroll
            for (int i = 0; i < 8; i++) hashes[(size_t)leaf * 8 + i] = bswap32(st[i]);   // digest bytes in memory order
        }
    }
}

// parents[i] = SHA-256(children[2i] || children[2i+1]); the last child of an odd level is paired with itself (tree.go:122-130)
__global__ void k_sha256_nodes(const uint32_t *__restrict__ children, uint32_t n_children, uint32_t *__restrict__ parents) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n_par = (n_children + 1) / 2;
    if (i >= n_par) return;
    const uint32_t l = 2 * i, r = (2 * i + 1 < n_children) ? 2 * i + 1 : l;
    uint32_t st[8], w[16];
    sha256_init(st);
#pragma unroll
    for (int j = 0; j < 8; j++) { w[j] = bswap32(children[(size_t)l * 8 + j]); w[8 + j] = bswap32(children[(size_t)r * 8 + j]); }
    sha256_compress(st, w);
    w[0] = 0x80000000u;
#pragma unroll
    for (int j = 1; j < 15; j++) w[j] = 0;
    w[15] = 512;
    sha256_compress(st, w);
#pragma unroll
    for (int j = 0; j < 8; j++) parents[(size_t)i * 8 + j] = bswap32(st[j]);
}

void launch_leaf_serialize(const Launch &L, const LeafFormat &F, const u64 *cts, uint64_t ct_words, uint32_t n_leaves, u64 *out) {
    LaunchScope scope(L, KID_LEAF_SER);
    const uint32_t words = F.padded_len / 8;
    const uint32_t bx = (words / 2 + 255) / 256;
    k_leaf_serialize<<<bx * n_leaves, 256, 0, L.stream>>>(F, cts, ct_words, out, bx);
}
void launch_ct_stream(const Launch &L, const LeafFormat &F, const u64 *cts, uint64_t ct_words, uint64_t byte0, uint64_t byte_end, u64 *out) {
    LaunchScope scope(L, KID_LEAF_SER);
    const uint64_t words = (byte_end - byte0 + 7) / 8;
    k_ct_stream<<<(unsigned)((words + 255) / 256), 256, 0, L.stream>>>(F, cts, ct_words, byte0, byte_end, out);
}
void launch_sha256_leaves(const Launch &L, const u64 *ser, uint32_t padded_len, uint32_t n_leaves, uint32_t *hashes) {
    LaunchScope scope(L, KID_SHA_LEAF);
    k_sha256_leaves<<<(n_leaves + 31) / 32, 32 * (1 + SHA_PRODUCERS), 0, L.stream>>>(reinterpret_cast<const uint4 *>(ser), padded_len / 64, n_leaves, hashes);
}
void launch_sha256_nodes(const Launch &L, const uint32_t *children, uint32_t n_children, uint32_t *parents) {
    LaunchScope scope(L, KID_SHA_NODE);
    const uint32_t np = (n_children + 1) / 2;
    k_sha256_nodes<<<(np + 127) / 128, 128, 0, L.stream>>>(children, n_children, parents);
}

}  // namespace lmn
