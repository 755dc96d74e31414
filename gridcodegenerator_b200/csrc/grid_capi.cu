// grid_capi.cu -- extern "C" surface (include/grid_b200.h) over one emitted grid.cuh.
// Compiled once per robot:  nvcc -gencode arch=compute_100a,code=sm_100a -I<dir with grid.cuh> ...
#include "grid.cuh"
#include "grid_b200.h"

#include <mutex>
#include <string>

#ifndef GRID_ROBOT_NAME
#define GRID_ROBOT_NAME "robot"
#endif

namespace {

thread_local std::string g_last_error;

int fail(cudaError_t e, const char* where) {
    if (e == cudaSuccess) return 0;
    g_last_error = std::string(where) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
    return static_cast<int>(e);
}
int bad_arg(const char* where, const char* what) {
    g_last_error = std::string(where) + ": " + what;
    return static_cast<int>(cudaErrorInvalidValue);
}

constexpr int MAX_DEVICES = 16;

// robot model handle (owns the optional global state scratch) per device and scalar type
template <typename T>
struct Models {
    grid::robotModel<T>* d_model[MAX_DEVICES] = {};
    std::mutex mu;
    grid::robotModel<T>* get() {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) return nullptr;
        std::lock_guard<std::mutex> lk(mu);
        if (!d_model[dev]) d_model[dev] = grid::init_robotModel<T>();
        return d_model[dev];
    }
    void release() {
        std::lock_guard<std::mutex> lk(mu);
        int cur = 0;
        cudaGetDevice(&cur);
        for (int d = 0; d < MAX_DEVICES; d++)
            if (d_model[d]) {
                cudaSetDevice(d);
                grid::close_grid<T>(nullptr, d_model[d], nullptr);
                d_model[d] = nullptr;
            }
        cudaSetDevice(cur);
    }
};
template <typename T>
Models<T>& models() { static Models<T> m; return m; }

// cached device workspace + streams of the host-buffer entry points (current device only)
struct HostWork {
    void* d_in = nullptr; void* d_in2 = nullptr; void* d_out = nullptr;
    size_t in_bytes = 0, in2_bytes = 0, out_bytes = 0;
    cudaStream_t streams[3] = {};
    cudaEvent_t ev[3] = {};
    bool have_streams = false;
    int device = -1;
    std::mutex mu;
    cudaError_t ensure(size_t in_b, size_t in2_b, size_t out_b) {
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return e;
        if (dev != device) { free_all(); device = dev; }
        if (!have_streams) {
            for (int i = 0; i < 3; i++) {
                if ((e = cudaStreamCreateWithFlags(&streams[i], cudaStreamNonBlocking)) != cudaSuccess) return e;
                if ((e = cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming)) != cudaSuccess) return e;
            }
            have_streams = true;
        }
        if ((e = grow(&d_in, &in_bytes, in_b)) != cudaSuccess) return e;
        if ((e = grow(&d_in2, &in2_bytes, in2_b)) != cudaSuccess) return e;
        return grow(&d_out, &out_bytes, out_b);
    }
    static cudaError_t grow(void** p, size_t* have, size_t want) {
        if (want <= *have) return cudaSuccess;
        if (*p) cudaFree(*p);
        *p = nullptr; *have = 0;
        cudaError_t e = cudaMalloc(p, want);
        if (e == cudaSuccess) *have = want;
        return e;
    }
    void free_all() {
        if (d_in) cudaFree(d_in);
        if (d_in2) cudaFree(d_in2);
        if (d_out) cudaFree(d_out);
        d_in = d_in2 = d_out = nullptr; in_bytes = in2_bytes = out_bytes = 0;
        if (have_streams) for (int i = 0; i < 3; i++) { cudaStreamDestroy(streams[i]); cudaEventDestroy(ev[i]); }
        have_streams = false;
    }
};
HostWork& host_work() { static HostWork w; return w; }

constexpr int N = grid::NUM_JOINTS;

// ---- device-pointer implementations
template <typename T>
int inverse_dynamics_impl(const T* d_q_qd, int stride, const T* d_qdd, T* d_c, int nk, T g, void* s) {
    const char* fn = "grid_inverse_dynamics";
    if (nk < 0 || stride < 2 * N || (nk > 0 && (!d_q_qd || !d_c))) return bad_arg(fn, "bad pointer / stride / num_knots");
    auto* m = models<T>().get();
    if (!m) return bad_arg(fn, "no CUDA device");
    cudaStream_t st = static_cast<cudaStream_t>(s);
    return fail(d_qdd ? grid::inverse_dynamics_qdd_launch<T>(d_c, d_q_qd, stride, d_qdd, m, g, nk, st)
                      : grid::inverse_dynamics_launch<T>(d_c, d_q_qd, stride, m, g, nk, st), fn);
}
template <typename T>
int direct_minv_impl(const T* d_q, int stride, T* d_Minv, int nk, void* s) {
    const char* fn = "grid_direct_minv";
    if (nk < 0 || stride < N || (nk > 0 && (!d_q || !d_Minv))) return bad_arg(fn, "bad pointer / stride / num_knots");
    auto* m = models<T>().get();
    if (!m) return bad_arg(fn, "no CUDA device");
    return fail(grid::direct_minv_launch<T>(d_Minv, d_q, stride, m, nk, static_cast<cudaStream_t>(s)), fn);
}
template <typename T>
int forward_dynamics_impl(const T* d_in, int stride, T* d_qdd, int nk, T g, void* s) {
    const char* fn = "grid_forward_dynamics";
    if (nk < 0 || stride < 3 * N || (nk > 0 && (!d_in || !d_qdd))) return bad_arg(fn, "bad pointer / stride / num_knots");
    auto* m = models<T>().get();
    if (!m) return bad_arg(fn, "no CUDA device");
    return fail(grid::forward_dynamics_launch<T>(d_qdd, d_in, stride, m, g, nk, static_cast<cudaStream_t>(s)), fn);
}
template <typename T>
int inverse_dynamics_gradient_impl(const T* d_q_qd, int stride, const T* d_qdd, T* d_out, int nk, T g, void* s) {
    const char* fn = "grid_inverse_dynamics_gradient";
    if (nk < 0 || stride < 2 * N || (nk > 0 && (!d_q_qd || !d_out))) return bad_arg(fn, "bad pointer / stride / num_knots");
    auto* m = models<T>().get();
    if (!m) return bad_arg(fn, "no CUDA device");
    cudaStream_t st = static_cast<cudaStream_t>(s);
    return fail(d_qdd ? grid::inverse_dynamics_gradient_qdd_launch<T>(d_out, d_q_qd, stride, d_qdd, m, g, nk, st)
                      : grid::inverse_dynamics_gradient_launch<T>(d_out, d_q_qd, stride, m, g, nk, st), fn);
}
template <typename T>
int forward_dynamics_gradient_impl(const T* d_in, int stride, T* d_out, int nk, T g, void* s) {
    const char* fn = "grid_forward_dynamics_gradient";
    if (nk < 0 || stride < 3 * N || (nk > 0 && (!d_in || !d_out))) return bad_arg(fn, "bad pointer / stride / num_knots");
    auto* m = models<T>().get();
    if (!m) return bad_arg(fn, "no CUDA device");
    return fail(grid::forward_dynamics_gradient_launch<T>(d_out, d_in, stride, m, g, nk, static_cast<cudaStream_t>(s)), fn);
}
template <typename T>
int forward_dynamics_gradient_qdd_minv_impl(const T* d_q_qd, int stride, const T* d_qdd, const T* d_Minv, T* d_out,
                                            int nk, T g, void* s) {
    const char* fn = "grid_forward_dynamics_gradient_qdd_minv";
    if (nk < 0 || stride < 2 * N || (nk > 0 && (!d_q_qd || !d_qdd || !d_Minv || !d_out)))
        return bad_arg(fn, "bad pointer / stride / num_knots");
    auto* m = models<T>().get();
    if (!m) return bad_arg(fn, "no CUDA device");
    return fail(grid::forward_dynamics_gradient_qdd_minv_launch<T>(d_out, d_q_qd, stride, d_qdd, d_Minv, m, g, nk,
                                                                   static_cast<cudaStream_t>(s)), fn);
}

template <typename T>
int aba_impl(const T* d_in, int stride, T* d_qdd, int nk, T g, void* s) {
    const char* fn = "grid_aba";
    if (nk < 0 || stride < 3 * N || (nk > 0 && (!d_in || !d_qdd))) return bad_arg(fn, "bad pointer / stride / num_knots");
    auto* m = models<T>().get();
    if (!m) return bad_arg(fn, "no CUDA device");
    return fail(grid::aba_launch<T>(d_qdd, d_in, stride, m, g, nk, static_cast<cudaStream_t>(s)), fn);
}
template <typename T>
int crba_impl(const T* d_q, int stride, T* d_M, int nk, void* s) {
    const char* fn = "grid_crba";
    if (nk < 0 || stride < N || (nk > 0 && (!d_q || !d_M))) return bad_arg(fn, "bad pointer / stride / num_knots");
    auto* m = models<T>().get();
    if (!m) return bad_arg(fn, "no CUDA device");
    return fail(grid::crba_launch<T>(d_M, d_q, stride, m, T(0), nk, static_cast<cudaStream_t>(s)), fn);
}

// ---- host-buffer pipeline: chunks of the batch rotate over three streams so that the H2D copy of
// chunk i+1, the kernel of chunk i and the D2H copy of chunk i-1 overlap (PCIe is full duplex).
template <typename T, class Launch>
int host_pipeline(const char* fn, const T* h_in, int in_per_knot, T* h_out, int out_per_knot, int nk, Launch launch) {
    if (nk < 0 || (nk > 0 && (!h_in || !h_out))) return bad_arg(fn, "bad pointer / num_knots");
    if (nk == 0) return 0;
    HostWork& w = host_work();
    std::lock_guard<std::mutex> lk(w.mu);
    cudaError_t e = w.ensure(size_t(nk) * in_per_knot * sizeof(T), 0, size_t(nk) * out_per_knot * sizeof(T));
    if (e != cudaSuccess) return fail(e, fn);
    const int chunk = 16384;                      // multiple of the 32-knot tile: every chunk stays TMA-aligned
    T* d_in = static_cast<T*>(w.d_in);
    T* d_out = static_cast<T*>(w.d_out);
    int i = 0;
    for (int k0 = 0; k0 < nk; k0 += chunk, i++) {
        const int c = nk - k0 < chunk ? nk - k0 : chunk;
        cudaStream_t st = w.streams[i % 3];
        e = cudaMemcpyAsync(d_in + size_t(k0) * in_per_knot, h_in + size_t(k0) * in_per_knot,
                            size_t(c) * in_per_knot * sizeof(T), cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) return fail(e, fn);
        int rc = launch(d_in + size_t(k0) * in_per_knot, d_out + size_t(k0) * out_per_knot, c, st);
        if (rc) return rc;
        e = cudaMemcpyAsync(h_out + size_t(k0) * out_per_knot, d_out + size_t(k0) * out_per_knot,
                            size_t(c) * out_per_knot * sizeof(T), cudaMemcpyDeviceToHost, st);
        if (e != cudaSuccess) return fail(e, fn);
    }
    for (int s = 0; s < 3; s++)
        if ((e = cudaStreamSynchronize(w.streams[s])) != cudaSuccess) return fail(e, fn);
    return 0;
}

}  // namespace

#define GRID_INFO_CASE(IDX, CFG, ALGO, KPTR)                                                                      \
    case IDX: {                                                                                                   \
        if (what == 0) return long(grid::CFG<T>::SMEM_BYTES);                                                     \
        if (what == 1) return long(grid::CFG<T>::NT);                                                             \
        if (what == 2) return long(grid::ALGO<T>::SLOTS);                                                         \
        if (what == 3) return long(grid::CFG<T>::STATE_SMEM);                                                     \
        if (what == 4) return long(grid::CFG<T>::STAGED);                                                         \
        if (what == 5) return long(grid_rt::persistent_blocks(reinterpret_cast<const void*>(KPTR),                \
                                                              grid::CFG<T>::NT, grid::CFG<T>::SMEM_BYTES, 1 << 20)); \
        return -1;                                                                                                \
    }

template <typename T>
long kernel_info(int algo, int what) {
    using namespace grid;
    using RM = const robotModel<T>*;
    switch (algo) {
        GRID_INFO_CASE(0, inverse_dynamics_cfg, inverse_dynamics_algo,
                       static_cast<void (*)(T*, const T*, const int, RM, const T, const int)>(inverse_dynamics_kernel<T>))
        GRID_INFO_CASE(1, inverse_dynamics_qdd_cfg, inverse_dynamics_qdd_algo,
                       static_cast<void (*)(T*, const T*, const int, const T*, RM, const T, const int)>(inverse_dynamics_kernel<T>))
        GRID_INFO_CASE(2, direct_minv_cfg, direct_minv_algo,
                       static_cast<void (*)(T*, const T*, const int, RM, const int)>(direct_minv_kernel<T>))
        GRID_INFO_CASE(3, forward_dynamics_cfg, forward_dynamics_algo,
                       static_cast<void (*)(T*, const T*, const int, RM, const T, const int)>(forward_dynamics_kernel<T>))
        GRID_INFO_CASE(4, inverse_dynamics_gradient_cfg, inverse_dynamics_gradient_algo,
                       static_cast<void (*)(T*, const T*, const int, RM, const T, const int)>(inverse_dynamics_gradient_kernel<T>))
        GRID_INFO_CASE(5, inverse_dynamics_gradient_qdd_cfg, inverse_dynamics_gradient_qdd_algo,
                       static_cast<void (*)(T*, const T*, const int, const T*, RM, const T, const int)>(inverse_dynamics_gradient_kernel<T>))
        GRID_INFO_CASE(6, forward_dynamics_gradient_cfg, forward_dynamics_gradient_algo,
                       static_cast<void (*)(T*, const T*, const int, RM, const T, const int)>(forward_dynamics_gradient_kernel<T>))
        GRID_INFO_CASE(7, forward_dynamics_gradient_qdd_minv_cfg, forward_dynamics_gradient_qdd_minv_algo,
                       static_cast<void (*)(T*, const T*, const int, const T*, const T*, RM, const T, const int)>(forward_dynamics_gradient_kernel<T>))
        GRID_INFO_CASE(8, aba_cfg, aba_algo,
                       static_cast<void (*)(T*, const T*, const int, RM, const T, const int)>(aba_kernel<T>))
        GRID_INFO_CASE(9, crba_cfg, crba_algo,
                       static_cast<void (*)(T*, const T*, const int, RM, const T, const int)>(crba_kernel<T>))
    }
    return -1;
}
extern "C" {

int grid_num_joints(void) { return N; }
const char* grid_robot_name(void) { return GRID_ROBOT_NAME; }
const char* grid_last_error(void) { return g_last_error.c_str(); }
int grid_abi_version(void) { return 1; }

long grid_kernel_info(int algo, int is_f64, int what) {
    return is_f64 ? kernel_info<double>(algo, what) : kernel_info<float>(algo, what);
}

#define GRID_DEFINE(SUF, T)                                                                                        \
    int grid_inverse_dynamics_##SUF(const T* a, int st, const T* b, T* o, int nk, T g, void* s) {                 \
        return inverse_dynamics_impl<T>(a, st, b, o, nk, g, s); }                                                  \
    int grid_direct_minv_##SUF(const T* a, int st, T* o, int nk, void* s) {                                        \
        return direct_minv_impl<T>(a, st, o, nk, s); }                                                             \
    int grid_forward_dynamics_##SUF(const T* a, int st, T* o, int nk, T g, void* s) {                              \
        return forward_dynamics_impl<T>(a, st, o, nk, g, s); }                                                     \
    int grid_inverse_dynamics_gradient_##SUF(const T* a, int st, const T* b, T* o, int nk, T g, void* s) {        \
        return inverse_dynamics_gradient_impl<T>(a, st, b, o, nk, g, s); }                                         \
    int grid_forward_dynamics_gradient_##SUF(const T* a, int st, T* o, int nk, T g, void* s) {                     \
        return forward_dynamics_gradient_impl<T>(a, st, o, nk, g, s); }                                            \
    int grid_forward_dynamics_gradient_qdd_minv_##SUF(const T* a, int st, const T* b, const T* c, T* o, int nk,   \
                                                      T g, void* s) {                                              \
        return forward_dynamics_gradient_qdd_minv_impl<T>(a, st, b, c, o, nk, g, s); }                             \
    int grid_aba_##SUF(const T* a, int st, T* o, int nk, T g, void* s) { return aba_impl<T>(a, st, o, nk, g, s); }  \
    int grid_crba_##SUF(const T* a, int st, T* o, int nk, void* s) { return crba_impl<T>(a, st, o, nk, s); }      \
    int grid_forward_dynamics_gradient_host_##SUF(const T* h_in, T* h_out, int nk, T g) {                          \
        return host_pipeline<T>("grid_forward_dynamics_gradient_host", h_in, 3 * N, h_out, 2 * N * N, nk,          \
            [g](const T* di, T* dout, int c, cudaStream_t st) {                                                    \
                return forward_dynamics_gradient_impl<T>(di, 3 * N, dout, c, g, st); }); }

GRID_DEFINE(f32, float)
GRID_DEFINE(f64, double)

int grid_forward_dynamics_host_f32(const float* h_in, float* h_out, int nk, float g) {
    return host_pipeline<float>("grid_forward_dynamics_host", h_in, 3 * N, h_out, N, nk,
        [g](const float* di, float* dout, int c, cudaStream_t st) {
            return forward_dynamics_impl<float>(di, 3 * N, dout, c, g, st); });
}
// RNEA over a [q|qd|qdd] host batch: qdd is the third block of each knot's row
int grid_inverse_dynamics_host_f32(const float* h_in, float* h_out, int nk, float g) {
    const char* fn = "grid_inverse_dynamics_host";
    if (nk < 0 || (nk > 0 && (!h_in || !h_out))) return bad_arg(fn, "bad pointer / num_knots");
    if (nk == 0) return 0;
    HostWork& w = host_work();
    std::lock_guard<std::mutex> lk(w.mu);
    cudaError_t e = w.ensure(size_t(nk) * 3 * N * sizeof(float), size_t(nk) * N * sizeof(float), size_t(nk) * N * sizeof(float));
    if (e != cudaSuccess) return fail(e, fn);
    cudaStream_t st = w.streams[0];
    float* d_in = static_cast<float*>(w.d_in);
    float* d_qdd = static_cast<float*>(w.d_in2);
    float* d_out = static_cast<float*>(w.d_out);
    if ((e = cudaMemcpyAsync(d_in, h_in, size_t(nk) * 3 * N * sizeof(float), cudaMemcpyHostToDevice, st)) != cudaSuccess) return fail(e, fn);
    if ((e = cudaMemcpy2DAsync(d_qdd, N * sizeof(float), d_in + 2 * N, 3 * N * sizeof(float), N * sizeof(float), nk,
                               cudaMemcpyDeviceToDevice, st)) != cudaSuccess) return fail(e, fn);
    int rc = inverse_dynamics_impl<float>(d_in, 3 * N, d_qdd, d_out, nk, g, st);
    if (rc) return rc;
    if ((e = cudaMemcpyAsync(h_out, d_out, size_t(nk) * N * sizeof(float), cudaMemcpyDeviceToHost, st)) != cudaSuccess) return fail(e, fn);
    return fail(cudaStreamSynchronize(st), fn);
}

int grid_release(void) {
    host_work().free_all();
    models<float>().release();
    models<double>().release();
    return 0;
}

}  // extern "C"
