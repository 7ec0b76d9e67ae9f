// Constants shared by the tensor-core MLP kernels (forward: mlp_tc.cu, backward: mlp_bwd.cu).
#pragma once
#include <stdint.h>

namespace nb {

constexpr int kTileM = 128;
constexpr int kSlabBytes = 16384;  // [128 x 64] bf16
constexpr int kStages = 3;
constexpr int kNumSlabs = 73;
constexpr int kConstFloats = 3200;
// Saved activations (training): per 128-sample tile, 40 slabs of 16 KB, each the same
// [128 x 64] bf16 K-major SWIZZLE_128B image the kernels use in shared memory, so the backward
// kernels can bulk-copy them straight into UMMA operands:
//   slabs 4l..4l+3 = h_l (l = 0..7) | 32..35 = feature | 36,37 = view hidden | 38 = enc(pts) | 39 = enc(dirs)
constexpr int kActSlabs = 40;
constexpr int kActSlabFeat = 32, kActSlabHv = 36, kActSlabEncP = 38, kActSlabEncD = 39;
// relu' bit masks (training): after ALL activation tiles of a launch, per tile 9 groups (h_0..h_7, view hidden) x
// 8 chunks x 128 rows of uint32: bit 31-j of word [group][chunk][row] = (pre-activation of column 32*chunk + j > 0).
// The dgrad chain reads these 288 B per sample instead of the 4.35 KB of bf16 activations they summarise.
constexpr int kMaskGroups = 9;
constexpr int kMaskTileBytes = kMaskGroups * 8 * 128 * 4;   // 36864
// Pre-activation gradients written by the dgrad chain (same slab image):
//   slabs 4l..4l+3 = d_a_l (l = 0..7) | 32..35 = d_feature | 36,37 = d_a(view hidden)
constexpr int kDaSlabs = 38;
constexpr int kDaSlabFeat = 32, kDaSlabHv = 36;

}  // namespace nb
