// Workspace addressing shared by host code and kernels (layout documented in erasurehead.h).
#pragma once
#include "common.cuh"

namespace eh {

__host__ __device__ __forceinline__ CtlWord* ws_ctl(void* ws, const eh_ws_layout& lay) {
    return reinterpret_cast<CtlWord*>(reinterpret_cast<char*>(ws) + lay.off_ctl);
}
__host__ __device__ __forceinline__ uint32_t* ws_beta_flag(void* ws, const eh_ws_layout& lay) {
    return &ws_ctl(ws, lay)[CTL_BETA_FLAG].v;
}
__host__ __device__ __forceinline__ uint32_t* ws_error(void* ws, const eh_ws_layout& lay) {
    return &ws_ctl(ws, lay)[CTL_ERROR].v;
}
__host__ __device__ __forceinline__ float* ws_beta(void* ws, const eh_ws_layout& lay, int parity) {
    return reinterpret_cast<float*>(reinterpret_cast<char*>(ws) + lay.off_beta) + (size_t)parity * lay.d_pad;
}
__host__ __device__ __forceinline__ uint32_t* ws_flag(void* ws, const eh_ws_layout& lay, int w, int m) {
    return reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(ws) + lay.off_flags) +
           ((size_t)w * EH_MAX_MSGS + m) * 32;
}
__host__ __device__ __forceinline__ float* ws_slot(void* ws, const eh_ws_layout& lay, int w, int m, int parity) {
    return reinterpret_cast<float*>(reinterpret_cast<char*>(ws) + lay.off_slots) +
           (((size_t)w * lay.n_msgs + m) * 2 + parity) * lay.d_pad;
}

}  // namespace eh
