// Shared device helpers: sortable (score,id) keys, smem bitonic sort, vector loads.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace brag {

// ---------------------------------------------------------------- sortable keys
// key = ordered(score) << 32 | (0xFFFFFFFF - row).  Bigger key == better under
// (score desc, row asc).  key 0 is the empty slot.
__device__ __forceinline__ uint32_t f32_ordered(float f) {
    f += 0.0f;  // -0 -> +0 so equal scores compare equal
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ordered_f32(uint32_t k) {
    uint32_t u = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
    return __uint_as_float(u);
}
__device__ __forceinline__ uint64_t make_key(float s, uint32_t row) {
    return ((uint64_t)f32_ordered(s) << 32) | (uint64_t)(0xFFFFFFFFu - row);
}
__device__ __forceinline__ float key_score(uint64_t k) { return ordered_f32((uint32_t)(k >> 32)); }
__device__ __forceinline__ uint32_t key_row(uint64_t k) { return 0xFFFFFFFFu - (uint32_t)k; }

// ---------------------------------------------------------------- bitonic sort (descending)
// n must be a power of two; all `nthreads` threads of the group call it; `sync` is
// __syncthreads for a CTA-wide sort or __syncwarp for a warp-private one.
template <typename T, typename SyncF>
__device__ __forceinline__ void bitonic_desc(T* a, int n, int tid, int nthreads, SyncF sync) {
    for (int k = 2; k <= n; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < (n >> 1); i += nthreads) {
                // i-th compare-exchange pair of this stage
                int lo = ((i & ~(j - 1)) << 1) | (i & (j - 1));
                int hi = lo | j;
                bool desc = (lo & k) == 0;
                T x = a[lo], y = a[hi];
                if ((x < y) == desc) { a[lo] = y; a[hi] = x; }
            }
            sync();
        }
    }
}

__device__ __forceinline__ int next_pow2(int x) {
    int p = 1;
    while (p < x) p <<= 1;
    return p;
}

// ---------------------------------------------------------------- loads
__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

__device__ __forceinline__ bool mask_test(const uint32_t* __restrict__ mask, uint32_t row) {
    return mask == nullptr || ((__ldg(mask + (row >> 5)) >> (row & 31)) & 1u);
}

// unpack 8 storage elements (one 16-byte chunk) to float
template <int DT> struct Unpack;
template <> struct Unpack<1> {  // f16
    __device__ static __forceinline__ void run(const uint4& v, float* f) {
        const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
        for (int i = 0; i < 4; ++i) { float2 t = __half22float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
    }
};
template <> struct Unpack<2> {  // bf16
    __device__ static __forceinline__ void run(const uint4& v, float* f) {
        const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
        for (int i = 0; i < 4; ++i) { f[2 * i] = __uint_as_float(u[i] << 16); f[2 * i + 1] = __uint_as_float(u[i] & 0xFFFF0000u); }
    }
};

}  // namespace brag
