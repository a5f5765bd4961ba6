// Second copy of the persistent loop kernel, compiled WITHOUT the arbitrary-length tier (RELP_NO_WCB) in its own
// namespace: instruction for instruction the int64 / short-multi-limb loop, with nothing of relp_wcb.cuh in its call
// graph.  Measured on the headline LP (B200, same box): 30.5 k pivots/s, against 24.4 k for the copy whose slow paths
// can reach the cooperative routines (register allocation around the calls and code layout; the hot instructions are
// the same).  The host (relp_capi.cu launch_persistent) runs this copy until the device reports that a value longer
// than RELP_LONG_TRIGGER limbs has been stored (DState::long_seen -> want_long), then continues with the full copy.
#define RELP_NS relp_fast
#define RELP_NO_WCB 1
#include "relp_kernels.cuh"
#include <cstring>

using namespace relp_fast;

extern "C" {
// cudaOccupancyMaxActiveBlocksPerMultiprocessor of the kernel (0 dynamic shared memory)
int relp_fast_loop_occupancy(int* per_sm) {
    return (int)cudaOccupancyMaxActiveBlocksPerMultiprocessor(per_sm, k_simplex_loop<false>, RELP_PTHREADS, 0);
}
int relp_fast_loop_smem_optin(int bytes) {
    cudaError_t e = cudaFuncSetAttribute(k_simplex_loop<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return (int)e;
    return (int)cudaFuncSetAttribute(k_simplex_loop<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}
// `dev` points at the caller's Dev (the same struct, compiled from the same header in the caller's namespace)
int relp_fast_loop_launch(int peers, const void* dev, size_t dev_bytes, int rule, int max_iters, int smem_words, int use_narrow,
                          int blocks, size_t smem_bytes, void* stream) {
    if (dev_bytes != sizeof(Dev)) return (int)cudaErrorInvalidValue;
    Dev D;
    memcpy(&D, dev, sizeof(Dev));
    D.allow_long = 0;
    void* args[] = { (void*)&D, (void*)&rule, (void*)&max_iters, (void*)&smem_words, (void*)&use_narrow };
    return (int)cudaLaunchCooperativeKernel(peers ? (void*)k_simplex_loop<true> : (void*)k_simplex_loop<false>, dim3(blocks),
                                            dim3(RELP_PTHREADS), args, smem_bytes, (cudaStream_t)stream);
}
}
