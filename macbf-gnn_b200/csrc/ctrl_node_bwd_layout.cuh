// Per-tile record (128 agents) shared by the kernels of the tensor-core controller-node backward
// (ctrl_node_tc.cu <DUMP>, ctrl_node_bwd_tc.cu, wgrad_tc_kernel<3> in cbf_bwd_tc.cu): three 96-KB stages of bf16
// hi/lo operand images in the MN-major SWIZZLE_128B layout of cbf_bwd_layout.cuh, then ReLU masks and per-agent
// auxiliaries.
#pragma once
#include <stdint.h>

namespace macbf {
namespace cnrec {
constexpr uint32_t BLK = 128 * 128;
constexpr uint32_t STAGE = 6 * BLK;
// stage 0 (dW3^T = A2^T dZ3)
constexpr uint32_t A2_HI = 0, A2_LO = 2 * BLK, Z3_HI = 4 * BLK, Z3_LO = 5 * BLK;
// stage 1 (dW2 = dZ2^T A1)
constexpr uint32_t Z2_HI = STAGE, Z2_LO = STAGE + 2 * BLK, A1_HI = STAGE + 4 * BLK, A1_LO = STAGE + 5 * BLK;
// stage 2 (dW1[:, :128]^T = pool^T dZ1)
constexpr uint32_t P_HI = 2 * STAGE, P_LO = 2 * STAGE + 2 * BLK, Z1_HI = 2 * STAGE + 4 * BLK, Z1_LO = 2 * STAGE + 5 * BLK;
constexpr uint32_t MASKS = 3 * STAGE;                  // uint32 [3][4 cgrp][128]: layer 1 (16 bits), layer 2 (32), layer 3 (16)
constexpr uint32_t AUX = MASKS + 3 * 4 * 128 * 4;      // float [128][12]: d_o[4], direct d(e_x, e_y, v_x, v_y), (e_x, e_y, v_x, v_y)
constexpr uint32_t BYTES = AUX + 128 * 12 * 4;         // 307,200
static_assert(BYTES % 1024 == 0, "tile records must keep 1024-B alignment");
}  // namespace cnrec
}  // namespace macbf
