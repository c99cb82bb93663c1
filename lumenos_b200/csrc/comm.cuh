// Launchers of kernels_comm.cu (layout kernels around the NCCL exchanges of api_comm.cu).
#pragma once
#include "kernels.cuh"

namespace lmn {
void launch_slices_to_whole(const Launch &L, const u64 *stage, u64 *whole, uint32_t G, uint32_t J, uint32_t PL, uint32_t ns);
void launch_whole_to_slices(const Launch &L, const u64 *whole, u64 *stage, uint32_t G, uint32_t J, uint32_t PL, uint32_t ns);
void launch_copy_rows(const Launch &L, const u64 *src, const uint32_t *idx, u64 *dst, uint32_t rows, uint64_t row_words);
}  // namespace lmn
