// Replay -> device: the sequence gather of rlpyt's UniformSequenceReplayBuffer.extract_batch
// (rlpyt/replays/sequence/n_step.py:68-106 over rlpyt/utils/misc.py:38-56 extract_sequences) on a ring buffer that lives in
// HBM, instead of a numpy gather on the host followed by a 31 MB synchronous host-to-device copy per update
// (dreamer_algo.py:166-169).  Pure byte movement, HBM-bound:
//     dst[j, i, :] = src[(t0[i] + t_offset + j) mod T_cap, b[i], :]      j < T_seq, i < batch, rows of row_bytes bytes
// (t_offset = -1 for prev-action / prev-reward sequences; the modulo reproduces both wrap cases of extract_sequences).
#include "drm_common.cuh"

namespace drm {

__global__ void __launch_bounds__(256) replay_gather_kernel(const drm_replay_gather_call c) {
    const long long row = blockIdx.x;  // j * batch + i
    const int j = (int)(row / c.batch), i = (int)(row - (long long)j * c.batch);
    long long t = ((long long)c.t0[i] + c.t_offset + j) % c.T_cap;
    if (t < 0) t += c.T_cap;
    const unsigned char* src = static_cast<const unsigned char*>(c.src) + ((size_t)t * c.B_env + c.b[i]) * c.row_bytes;
    unsigned char* dst = static_cast<unsigned char*>(c.dst) + (size_t)row * c.row_bytes;
    if ((c.row_bytes & 15) == 0 && (((uintptr_t)c.src | (uintptr_t)c.dst) & 15) == 0) {
        const int n16 = (int)(c.row_bytes >> 4);
        const uint4* s4 = reinterpret_cast<const uint4*>(src);
        uint4* d4 = reinterpret_cast<uint4*>(dst);
        for (int k = threadIdx.x; k < n16; k += 256) d4[k] = __ldg(s4 + k);
    } else {
        for (long long k = threadIdx.x; k < c.row_bytes; k += 256) dst[k] = src[k];
    }
}

void warm_replay() { warm_kernel(replay_gather_kernel); }

}  // namespace drm

extern "C" int drm_replay_gather(const drm_replay_gather_call* c, void* stream_) {
    using namespace drm;
    cudaStream_t stream = (cudaStream_t)stream_;
    DRM_REQUIRE(c != nullptr, DRM_E_ARG, "null call");
    DRM_REQUIRE(c->T_cap >= 1 && c->B_env >= 1 && c->T_seq >= 1 && c->batch >= 1 && c->row_bytes >= 1, DRM_E_SHAPE, "bad dims");
    DRM_REQUIRE(c->src && c->dst && c->t0 && c->b, DRM_E_ARG, "null pointer");
    int rc = check_device(); if (rc) return rc;
    ProfScope prof(PROF_PACK, stream);
    replay_gather_kernel<<<(unsigned)((long long)c->T_seq * c->batch), 256, 0, stream>>>(*c);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}
