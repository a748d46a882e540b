// Allocation-free ModUp base conversion: b200fhe_bconv_modup_static.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  Same two kernels as
// b200fhe_bconv_modup (scale y_j = x_j (P/p_j)^-1 mod p_j, then the contraction), but the [m][N] scratch that holds y between
// them is NOT the plan's cudaMalloc'ed buffer: it is the file-scope device array below, B200FHE_STATIC_SCRATCH_WORDS
// 64-bit words (4 MiB), resident on every GPU that loads the library until the process exits.
//
// In host code `g_static_scratch` is only a host-side stand-in for the symbol, not a device address: the address is taken
// once per device with cudaGetSymbolAddress (that device current) and that pointer is what the kernels receive.
// Nothing has to start at zero: the scale kernel overwrites the m * N words the contraction then reads.  Every call on a
// device shares the one array, so calls that overlap on the GPU (different streams, or different plans) race on it;
// serialise them on one stream.  The entry point calls no cudaMalloc / cudaFree / allocating library routine.
#include "../../include/b200fhe.h"
#include "ctx.hpp"
#include <mutex>

int b200_set_err(int code, const char* fmt, ...);
void b200_count_launches(unsigned n);

static __device__ __align__(256) unsigned long long g_static_scratch[B200FHE_STATIC_SCRATCH_WORDS];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_mu;
uint64_t* g_scratch_ptr[kMaxDevices] = {};

// device address of the static scratch on `device` (which must be current); looked up once per device
int scratch_address(int device, uint64_t** out) {
    if (device < 0 || device >= kMaxDevices) return b200_set_err(B200FHE_EUNSUPPORTED, "bconv_modup_static: device index %d out of range", device);
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_ptr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p)
            return b200_set_err(B200FHE_ECUDA, "bconv_modup_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
        g_scratch_ptr[device] = (uint64_t*)p;
    }
    *out = g_scratch_ptr[device];
    return 0;
}
}  // namespace

extern "C" int b200fhe_bconv_modup_static(b200fhe_bconv* p, const uint64_t* d_in, uint64_t* d_out, void* stream) {
    if (!p || !d_in || !d_out) return b200_set_err(B200FHE_EINVAL, "bconv_modup_static: null argument");
    b200fhe_ctx* c = p->ctx;
    const size_t words = (size_t)p->m * c->N;
    if (words > (size_t)B200FHE_STATIC_SCRATCH_WORDS)
        return b200_set_err(B200FHE_EUNSUPPORTED,
                            "bconv_modup_static: m * N = %d * %zu = %zu words exceeds the static scratch of %zu words (%zu MiB); use b200fhe_bconv_modup",
                            p->m, c->N, words, (size_t)B200FHE_STATIC_SCRATCH_WORDS, (size_t)B200FHE_STATIC_SCRATCH_WORDS * 8 >> 20);
    if (p->mode != B200FHE_BCONV_HPS) return b200_set_err(B200FHE_EUNSUPPORTED, "bconv_modup_static: fast (HPS) conversion plans only");
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) return b200_set_err(B200FHE_ECUDA, "bconv_modup_static: %s", cudaGetErrorString(e));
    uint64_t* y = nullptr;
    if (int rc = scratch_address(c->device, &y)) return rc;
    if (int rc = b200fhe_bconv_scale(p, d_in, y, 0, (size_t)p->m, stream)) return rc;
    return b200fhe_bconv_contract(p, y, d_out, 0, (size_t)p->K, stream);
}
