// peer_dev.cuh — device-side flag primitives of the peer-memory frame (peer.cu), shared with k_finalize (post.cu),
// which carries the frame's last cross-GPU signals in its own prologue / tail instead of separate launches.
#pragma once
#include <stdint.h>

__device__ __forceinline__ uint32_t vx_ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void vx_st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
// spin until *p >= value (wrap-safe); ~4 s bound, then flag the error word and carry on (never hang the GPU)
__device__ __forceinline__ bool vx_spin_until(const uint32_t* p, uint32_t value, uint32_t* err) {
    if ((int32_t)(vx_ld_acquire_sys(p) - value) >= 0) return true;
    const long long t0 = clock64();
    while ((int32_t)(vx_ld_acquire_sys(p) - value) < 0) {
        if (clock64() - t0 > 8000000000LL) { atomicExch(err, value ? value : 1u); return false; }
        __nanosleep(32);
    }
    return true;
}
