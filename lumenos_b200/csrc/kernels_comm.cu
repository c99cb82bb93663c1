// Layout kernels of the multi-GPU exchanges: the encoded matrix exists as coefficient slices
// [ciphertext][poly][limb][N/G] during Encode and as whole ciphertexts [ciphertext][poly][limb][N] everywhere else.  NCCL moves
// contiguous buffers; these kernels turn G received slice blocks into whole ciphertexts (k_slices_to_whole), whole ciphertexts
// into G slice blocks to send (k_whole_to_slices), and gather selected ciphertext slices into a send buffer (k_copy_rows).
// HBM-bound copies: one 16-byte load and store per thread, consecutive threads on consecutive words of a run of N/G coefficients.
#include "comm.cuh"

namespace lmn {

// whole[j][pl][s * ns + c] = stage[s][j][pl][c]      (J ciphertexts, PL = 2 (level + 1) limbs, ns = N / G words per run)
__global__ void __launch_bounds__(256) k_slices_to_whole(const ulonglong2 *__restrict__ stage, ulonglong2 *__restrict__ whole, uint32_t G, uint32_t J,
                                                         uint32_t PL, uint32_t ns2 /* ns / 2 */) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t total = (uint64_t)G * J * PL * ns2;
    if (i >= total) return;
    const uint32_t c = (uint32_t)(i % ns2);
    uint64_t r = i / ns2;
    const uint32_t pl = (uint32_t)(r % PL); r /= PL;
    const uint32_t j = (uint32_t)(r % J), s = (uint32_t)(r / J);
    whole[(((uint64_t)j * PL + pl) * G + s) * ns2 + c] = stage[i];
}
// stage[s][j][pl][c] = whole[j][pl][s * ns + c]
__global__ void __launch_bounds__(256) k_whole_to_slices(const ulonglong2 *__restrict__ whole, ulonglong2 *__restrict__ stage, uint32_t G, uint32_t J,
                                                         uint32_t PL, uint32_t ns2) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t total = (uint64_t)G * J * PL * ns2;
    if (i >= total) return;
    const uint32_t c = (uint32_t)(i % ns2);
    uint64_t r = i / ns2;
    const uint32_t pl = (uint32_t)(r % PL); r /= PL;
    const uint32_t j = (uint32_t)(r % J), s = (uint32_t)(r / J);
    stage[i] = whole[(((uint64_t)j * PL + pl) * G + s) * ns2 + c];
}
// dst[r] = src[idx[r]] for rows of row_words words
__global__ void __launch_bounds__(256) k_copy_rows(const ulonglong2 *__restrict__ src, const uint32_t *__restrict__ idx, ulonglong2 *__restrict__ dst,
                                                   uint32_t rows, uint64_t row2 /* row_words / 2 */) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * row2) return;
    const uint32_t r = (uint32_t)(i / row2);
    dst[i] = src[(uint64_t)idx[r] * row2 + (i - (uint64_t)r * row2)];
}

void launch_slices_to_whole(const Launch &L, const u64 *stage, u64 *whole, uint32_t G, uint32_t J, uint32_t PL, uint32_t ns) {
    LaunchScope scope(L, KID_PACK);
    const uint64_t total = (uint64_t)G * J * PL * (ns / 2);
    k_slices_to_whole<<<(unsigned)((total + 255) / 256), 256, 0, L.stream>>>(reinterpret_cast<const ulonglong2 *>(stage), reinterpret_cast<ulonglong2 *>(whole), G, J, PL, ns / 2);
}
void launch_whole_to_slices(const Launch &L, const u64 *whole, u64 *stage, uint32_t G, uint32_t J, uint32_t PL, uint32_t ns) {
    LaunchScope scope(L, KID_PACK);
    const uint64_t total = (uint64_t)G * J * PL * (ns / 2);
    k_whole_to_slices<<<(unsigned)((total + 255) / 256), 256, 0, L.stream>>>(reinterpret_cast<const ulonglong2 *>(whole), reinterpret_cast<ulonglong2 *>(stage), G, J, PL, ns / 2);
}
void launch_copy_rows(const Launch &L, const u64 *src, const uint32_t *idx, u64 *dst, uint32_t rows, uint64_t row_words) {
    LaunchScope scope(L, KID_GATHER);
    const uint64_t total = rows * (row_words / 2);
    k_copy_rows<<<(unsigned)((total + 255) / 256), 256, 0, L.stream>>>(reinterpret_cast<const ulonglong2 *>(src), idx, reinterpret_cast<ulonglong2 *>(dst), rows, row_words / 2);
}

}  // namespace lmn
