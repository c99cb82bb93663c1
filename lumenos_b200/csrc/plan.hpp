// Encode plan: the reference's homomorphic "NTT" over the ciphertext index (fhe/ntt.go:20-281) compiled
// into a flat hypercube butterfly network.
//
// The reference mutates a slice of ciphertext pointers (core.Transpose, swaps) and applies ct+ct, ct-ct and
// ct*scalar in place.  Because every operation is an exact ring operation per (poly, limb, coeff) lane, the
// whole transform is a fixed linear map per lane.  The compiler below walks nttInner's control flow on
// *slot ids* (slot = index of the ciphertext in the input slice; data never moves) and records
//   - for every butterfly level s: the index bit the level pairs on (all pairs of a level differ in one bit),
//   - for every slot and level: the scalar (if any) multiplied into the slot after that level,
//   - the final position of every slot in the returned slice.
// Levels 0..split-1 touch only the high index bits (the size-n1 column transforms of the six-step split,
// fhe/ntt.go:254-258), the rest only the low bits, which gives the two HBM passes of the device kernel.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace lmn {

struct EncodePlan {
    uint32_t n = 0, logn = 0;
    uint32_t field_n = 0;                 // Field().N() the step arithmetic runs modulo (fhe/ntt.go:249,263)
    std::vector<uint8_t> stage_bit;       // [logn] bit paired by level s
    std::vector<uint16_t> scale_after;    // [logn][n] scalar id multiplied into slot after level s (0 = none)
    std::vector<uint32_t> out_pos;        // [n] position of slot in the returned slice
    std::vector<uint64_t> scalars;        // id -> plaintext-field scalar w (id 0 unused)
    uint32_t split = 0;                   // levels in pass 1 (0: single pass)
    uint32_t n1 = 0, n2 = 0;              // six-step factors at top level (n2 = stride of pass-1 tiles)
    uint64_t adds = 0, subs = 0, muls = 0;
};

// roots_forward: Field().RootForwardUint64(i), i < field_n.  t: plaintext modulus (for Pow(3, W(8))).
// max_tile: largest number of slots one device tile may hold.  Returns "" on success, else an error message.
std::string compile_encode_plan(uint32_t n, uint32_t field_n, const uint64_t *roots_forward, uint64_t t,
                                uint32_t max_tile, EncodePlan &plan);

}  // namespace lmn
