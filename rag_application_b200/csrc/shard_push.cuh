// Device side of ShardSink (kernels.h): a tail kernel that has ranked a query's final list stores it straight into the
// sharded call's destinations -- the step shard_finalize_push_kernel does for lists that still need a sort.
#pragma once
#include "common.cuh"
#include "kernels.h"

namespace brag {

// key (local row) of rank r < k of query b; key == 0 pads a short list
__device__ __forceinline__ void shard_sink_store(const ShardSink& sk, int B, int b, int k, int r, uint64_t key, int64_t id_offset) {
    if (key != 0ull)
        key = (key & 0xFFFFFFFF00000000ull) | (uint64_t)(0xFFFFFFFFu - (uint32_t)((int64_t)key_row(key) + id_offset));
    const size_t at = ((size_t)sk.list * B + b) * k + r;
    if (sk.local_out != nullptr) {
        sk.local_out[at] = key;
        return;
    }
    const unsigned ep = *sk.push.epoch;
    const size_t slot = sk.push.slots_off + ((size_t)(ep & 1u) * sk.push.world + sk.push.rank) * sk.push.slot_bytes;
    for (int g = 0; g < sk.push.world; ++g) reinterpret_cast<uint64_t*>(sk.push.windows[g] + slot)[at] = key;
}

}  // namespace brag
