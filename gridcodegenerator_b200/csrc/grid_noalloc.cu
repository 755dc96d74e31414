// grid_noalloc.cu -- ONE extern "C" entry point that allocates nothing (include/grid_b200.h:
// grid_forward_dynamics_gradient_noalloc_f32).  No counterpart in the reference: added
// for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
// that forbid it).
//
// The regular entry point grid_forward_dynamics_gradient_f32 takes its global scratch (per-thread state of robots
// whose state exceeds shared memory -- Atlas; lane-interleaved output rows of the transposed-store mode -- HyQ) from a
// robotModel handle that is cudaMalloc'ed on first use.  Here the same one-lane-per-knot tile loop runs with that
// scratch in a file-scope static __device__ array of fixed size: every GPU that loads this library holds
// GRID_NOALLOC_SCRATCH_BYTES of device memory until the process exits.
//
//  * limit: num_knots <= GRID_NOALLOC_MAX_KNOTS (8192; at most GRID_NOALLOC_MAX_BLOCKS = 32 CTAs run, the scratch holds 32 CTA slabs);
//    larger batches are rejected with cudaErrorInvalidValue and a message in grid_last_error();
//  * the array's name is a host-side stand-in in host code, not a device address: its device address is taken once per
//    device, with that device current, by cudaGetSymbolAddress and handed to the kernel as an argument;
//  * every call on a device shares the one array and finds the previous call's contents in it.  Nothing has to start at
//    zero: every state slot and every output-scratch element is written by a launch before that launch reads it.  Calls
//    that OVERLAP on the GPU (two streams, two host threads on one device) race on it -- serialise them;
//  * no cudaMalloc / cudaFree / allocating library routine is called, at set-up or per call.
#include "grid.cuh"
#include "grid_b200.h"

#include <mutex>
#include <string>

#define HIDDEN __attribute__((visibility("hidden")))
HIDDEN std::string& grid_capi_last_error();     // defined by the core part of grid_capi.cu (thread-local message)

namespace {

typedef grid::forward_dynamics_gradient_cfg<float> Cfg;
typedef grid::forward_dynamics_gradient_algo<float> Algo;

constexpr int N = grid::NUM_JOINTS;
constexpr int MAX_DEVICES = 16;
constexpr int NOALLOC_MAX_KNOTS = GRID_NOALLOC_MAX_KNOTS;
// CTAs that hold NOALLOC_MAX_KNOTS in one round of the tile loop, at most 32
constexpr int NOALLOC_TILES = (NOALLOC_MAX_KNOTS + grid_rt::TILE - 1) / grid_rt::TILE;
constexpr int NOALLOC_ROUND_BLOCKS = (NOALLOC_TILES + Cfg::NT / 32 - 1) / (Cfg::NT / 32);
constexpr int NOALLOC_MAX_BLOCKS = NOALLOC_ROUND_BLOCKS < 32 ? NOALLOC_ROUND_BLOCKS : 32;
// scratch elements one CTA of the tile loop addresses (grid_b200_runtime.cuh: tile_loop, S = a.scratch + blockIdx.x * ...)
constexpr size_t PER_BLOCK = (Cfg::STATE_SMEM ? 0 : size_t(Cfg::NT) * Algo::SLOTS) +
                             (Cfg::STAGED == 3 ? size_t(Cfg::NT) * Algo::OUT : 0);
constexpr size_t SCRATCH_ELEMS = PER_BLOCK * NOALLOC_MAX_BLOCKS > 64 ? PER_BLOCK * NOALLOC_MAX_BLOCKS : 64;
static_assert(SCRATCH_ELEMS * sizeof(float) <= (size_t(64) << 20), "no-alloc scratch is meant to stay modest (<= 64 MiB)");

}  // namespace

static __device__ __align__(256) float grid_noalloc_scratch[SCRATCH_ELEMS];

namespace {

__global__ void __maxnreg__((Cfg::NT * 255 <= 65536) ? 255 : (65536 / Cfg::NT / 8 * 8))
forward_dynamics_gradient_noalloc_kernel(float* d_df_du, const float* d_q_qd_u, const int stride, float* scratch,
                                         const float gravity, const int num_knots) {
    extern __shared__ __align__(16) unsigned char grid_smem[];
    grid_rt::TileArgs<float> a;
    a.in[0] = d_q_qd_u; a.stride[0] = stride;
    a.in[1] = nullptr; a.stride[1] = 0;
    a.in[2] = nullptr; a.stride[2] = 0;
    a.out = d_df_du; a.scratch = scratch;
    a.gravity = gravity; a.num_knots = num_knots;
    grid::forward_dynamics_gradient_device<float>(a, grid_smem);
}

struct Setup { float* scratch = nullptr; bool ready = false; };
Setup g_setup[MAX_DEVICES];
std::mutex g_mu;

int fail(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return 0;
    grid_capi_last_error() = std::string("grid_forward_dynamics_gradient_noalloc: ") + what + ": " + cudaGetErrorName(e) +
                             " (" + cudaGetErrorString(e) + ")";
    return static_cast<int>(e);
}

}  // namespace

extern "C" int grid_noalloc_max_knots(void) { return NOALLOC_MAX_KNOTS; }
extern "C" long grid_noalloc_scratch_bytes(void) { return long(SCRATCH_ELEMS * sizeof(float)); }

extern "C" int grid_forward_dynamics_gradient_noalloc_f32(const float* d_q_qd_u, int stride, float* d_df_du,
                                                          int num_knots, float gravity, void* stream) {
    if (num_knots < 0 || stride < 3 * N || (num_knots > 0 && (!d_q_qd_u || !d_df_du))) {
        grid_capi_last_error() = "grid_forward_dynamics_gradient_noalloc: bad pointer / stride / num_knots";
        return static_cast<int>(cudaErrorInvalidValue);
    }
    if (num_knots > NOALLOC_MAX_KNOTS) {
        grid_capi_last_error() = "grid_forward_dynamics_gradient_noalloc: num_knots " + std::to_string(num_knots) +
                                 " exceeds GRID_NOALLOC_MAX_KNOTS = " + std::to_string(NOALLOC_MAX_KNOTS) +
                                 " (the static scratch is sized for that; use grid_forward_dynamics_gradient_f32)";
        return static_cast<int>(cudaErrorInvalidValue);
    }
    if (num_knots == 0) return 0;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(e, "cudaGetDevice");
    if (dev < 0 || dev >= MAX_DEVICES) return fail(cudaErrorInvalidDevice, "device index");
    float* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        Setup& s = g_setup[dev];
        if (!s.ready) {
            // once per device, with that device current: the symbol's DEVICE address (the bare name is a host shadow)
            void* p = nullptr;
            if ((e = cudaGetSymbolAddress(&p, grid_noalloc_scratch)) != cudaSuccess) return fail(e, "cudaGetSymbolAddress");
            if (!p) return fail(cudaErrorInvalidSymbol, "cudaGetSymbolAddress returned a null address");
            if ((e = cudaFuncSetAttribute(forward_dynamics_gradient_noalloc_kernel,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg::SMEM_BYTES))) != cudaSuccess)
                return fail(e, "cudaFuncSetAttribute");
            s.scratch = static_cast<float*>(p);
            s.ready = true;
        }
        scratch = s.scratch;
    }
    const int tiles = (num_knots + grid_rt::TILE - 1) / grid_rt::TILE;
    const int blocks = tiles < NOALLOC_MAX_BLOCKS ? tiles : NOALLOC_MAX_BLOCKS;   // scratch is indexed by CTA
    forward_dynamics_gradient_noalloc_kernel<<<blocks, Cfg::NT, Cfg::SMEM_BYTES, static_cast<cudaStream_t>(stream)>>>(
        d_df_du, d_q_qd_u, stride, scratch, gravity, num_knots);
    return fail(cudaGetLastError(), "kernel launch");
}
