// Per-tile record shared by the three kernels of the tensor-core CBF backward (cbf_bwd_tc.cu):
//   K2b-1  cbf_forward_tc_kernel<true>   recompute + dump                      (writes A2, masks, xrec, gm)
//   K2b-2  cbf_dgrad_tc_kernel           delta chain on the tensor core       (writes dZ2)
//   K2b-3  wgrad_tc_kernel<2>            dW3, dW2, db3, db2 = GEMMs over edges (reads A2, dZ2; rebuilds a1, dZ3)
// A tile is 128 edge rows.  Activations / deltas are stored as bf16 (hi, lo) operand images in exactly
// the shared-memory layout the wgrad MMAs read (MN-major SWIZZLE_128B: per 64-channel block 128 rows x
// 128 B, 16-byte chunk c of row e at position c ^ (e & 7)), so K2b-3 stages them with plain bulk copies.
// The a1 and dZ3 images exist only in the stage buffers: the wgrad kernel rebuilds them on chip from XREC, GM and the
// layer-3 ReLU bits.
#pragma once
#include <stdint.h>

namespace macbf {
namespace cbfrec {
constexpr uint32_t BLK = 128 * 128;            // one 64-channel block of one image half: 16 KB
// In SHARED memory a stage is 96 KB: [A hi: 2 blocks][A lo: 2 blocks][B hi][B lo].  The record carries only the A parts
// (the B parts - dZ3 for stage G3, a1 for stage G2 - are rebuilt on chip), 64 KB per stage:
constexpr uint32_t STAGE_BYTES = 6 * BLK;      // a stage buffer in shared memory: 98,304
constexpr uint32_t REC_STAGE = 4 * BLK;        // bytes of a stage kept in the record: 65,536
constexpr uint32_t A2_HI = 0, A2_LO = 2 * BLK;                            // stage G3 (dW3^T = A2^T dZ3)
constexpr uint32_t Z2_HI = REC_STAGE, Z2_LO = REC_STAGE + 2 * BLK;        // stage G2 (dW2 = dZ2^T a1)
// per-thread ReLU masks (bit q = channel c0 + q was active), [which][cgrp][row] uint32
constexpr uint32_t MASKS = 2 * REC_STAGE;      // which: 0 = layer 1, 1/2 = layer 2 chunk 0/1, 3 = layer 3
constexpr uint32_t XREC = MASKS + 4 * 4 * 128 * 4;    // float[128][8]: the 6 edge features, d, unused
constexpr uint32_t GM = XREC + 128 * 8 * 4;           // float[128]: dL/dh * radius mask (0 for padding rows)
constexpr uint32_t GMASK = GM + 128 * 4;              // float[128]: the radius mask alone (recording forward: dL/dh comes later;
                                                      // cbf_dgrad_tc_kernel then writes gm = dL/dh * mask into GM)
constexpr uint32_t BYTES = GMASK + 128 * 4;           // 144,384 = 141 * 1024
constexpr uint32_t A3_BYTES = 4 * 128 * 16 * 4;       // recording forward only: layer-3 activations [cgrp][4][row][4] fp32, kept in a
                                                      // separate array behind the records (32 KB per tile)
static_assert(BYTES % 1024 == 0, "tile records must keep 1024-B alignment");
}  // namespace cbfrec
}  // namespace macbf
