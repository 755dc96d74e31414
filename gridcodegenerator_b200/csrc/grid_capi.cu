// grid_capi.cu -- extern "C" surface (include/grid_b200.h) over one emitted grid.cuh.
// Compiled per robot:  nvcc -gencode arch=compute_100a,code=sm_100a -I<dir with grid.cuh> ...
//
// The emitted kernels are large straight-line functions, so the build (runtime.build_library) compiles this file
// several times in parallel, each time with -DGRID_PARTS=<bitmask of algorithms> -DGRID_DTYPES=<1 f32 | 2 f64>,
// and links the objects into one libgrid_<robot>.so.  Compiled without those macros it is the whole library in
// one translation unit.  Shared state (robot-model handles, host workspace, last error) is defined by the part
// that has PART_CORE and dtype f32, and is referenced by the others through hidden-visibility symbols.
#include "grid.cuh"
#include "grid_b200.h"

#include <cstdlib>
#include <mutex>
#include <string>

#ifndef GRID_ROBOT_NAME
#define GRID_ROBOT_NAME "robot"
#endif
#define PART_CORE 0x01       // introspection, release, shared state, ID, MINV, FD
#define PART_ID_DU 0x02
#define PART_FD_DU 0x04
#define PART_FD_DU_QM 0x08
#define PART_ABA_CRBA 0x10
#define PART_IDSVA 0x20
#define PART_FDSVA 0x40
#define PART_EE 0x80
#define PART_FD_LIN 0x100
#ifndef GRID_PARTS
#define GRID_PARTS 0x1FF
#endif
#ifndef GRID_DTYPES
#define GRID_DTYPES 3
#endif
#define HAS(p) ((GRID_PARTS & (p)) != 0)
#define HIDDEN __attribute__((visibility("hidden")))

constexpr int MAX_DEVICES = 16;
constexpr int N = grid::NUM_JOINTS;

// Robot model handles per device and scalar type.  A handle owns the global scratch that some kernels use for
// per-thread state / transposed output (cfg::NEEDS_SCRATCH): the scratch is indexed by CTA, so at most ONE such
// kernel may be in flight per handle.  Launches are stream-ordered, hence one scratch-owning handle per
// (device, stream), created on the first scratch-using launch on that stream; kernels that never touch the
// scratch share one scratch-less handle per device.  (Round-1 had one scratch per device: two chunks of the host
// pipeline, or two user streams, ran Atlas / HyQ gradient kernels concurrently on the same scratch blocks.)
template <typename T>
struct Models {
    struct Owned { int dev; cudaStream_t stream; grid::robotModel<T>* m; unsigned long last_use; };
    static constexpr int MAX_OWNED = 32;
    grid::robotModel<T>* plain[MAX_DEVICES] = {};
    Owned owned[MAX_OWNED] = {};
    int nowned = 0;
    unsigned long tick = 0;
    std::mutex mu;
    grid::robotModel<T>* get(cudaStream_t st, bool needs_scratch) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) return nullptr;
        std::lock_guard<std::mutex> lk(mu);
        if (!needs_scratch) {
            if (!plain[dev]) plain[dev] = grid::init_robotModel<T>(false);
            return plain[dev];
        }
        ++tick;
        for (int i = 0; i < nowned; i++)
            if (owned[i].dev == dev && owned[i].stream == st) { owned[i].last_use = tick; return owned[i].m; }
        if (nowned < MAX_OWNED) {
            owned[nowned] = Owned{dev, st, grid::init_robotModel<T>(true), tick};
            return owned[nowned++].m;
        }
        // pool exhausted: hand the least recently used handle of this device to the new stream once its work is done
        int lru = -1;
        for (int i = 0; i < nowned; i++)
            if (owned[i].dev == dev && (lru < 0 || owned[i].last_use < owned[lru].last_use)) lru = i;
        if (lru < 0) return nullptr;
        if (cudaStreamSynchronize(owned[lru].stream) != cudaSuccess) { cudaGetLastError(); cudaDeviceSynchronize(); }
        owned[lru].stream = st; owned[lru].last_use = tick;
        return owned[lru].m;
    }
    void release() {
        std::lock_guard<std::mutex> lk(mu);
        int cur = 0;
        cudaGetDevice(&cur);
        for (int d = 0; d < MAX_DEVICES; d++)
            if (plain[d]) {
                cudaSetDevice(d);
                grid::close_grid<T>(nullptr, plain[d], nullptr);
                plain[d] = nullptr;
            }
        for (int i = 0; i < nowned; i++) {
            cudaSetDevice(owned[i].dev);
            grid::close_grid<T>(nullptr, owned[i].m, nullptr);
        }
        nowned = 0;
        cudaSetDevice(cur);
    }
};

// cached device workspace + streams of the host-buffer entry points, one per device (a host thread driving
// device d only ever takes device d's lock)
struct HostWork {
    void* d_in = nullptr; void* d_in2 = nullptr; void* d_out = nullptr;
    size_t in_bytes = 0, in2_bytes = 0, out_bytes = 0;
    cudaStream_t streams[3] = {};
    bool have_streams = false;
    std::mutex mu;
    cudaError_t ensure(size_t in_b, size_t in2_b, size_t out_b) {
        cudaError_t e;
        if (!have_streams) {
            for (int i = 0; i < 3; i++)
                if ((e = cudaStreamCreateWithFlags(&streams[i], cudaStreamNonBlocking)) != cudaSuccess) return e;
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
        if (have_streams) for (int i = 0; i < 3; i++) cudaStreamDestroy(streams[i]);
        have_streams = false;
    }
};

// ---- state shared by all parts of one library (hidden: not visible to other libgrid_<robot>.so files)
HIDDEN extern Models<float> grid_capi_models_f32;
HIDDEN extern Models<double> grid_capi_models_f64;
HIDDEN extern HostWork grid_capi_host_work[MAX_DEVICES];
HIDDEN std::string& grid_capi_last_error();
#if HAS(PART_CORE) && (GRID_DTYPES & 1)
Models<float> grid_capi_models_f32;
Models<double> grid_capi_models_f64;
HostWork grid_capi_host_work[MAX_DEVICES];
std::string& grid_capi_last_error() { static thread_local std::string s; return s; }
#endif

namespace {

template <typename T> Models<T>& models();
template <> Models<float>& models<float>() { return grid_capi_models_f32; }
template <> Models<double>& models<double>() { return grid_capi_models_f64; }

int fail(cudaError_t e, const char* where) {
    if (e == cudaSuccess) return 0;
    grid_capi_last_error() = std::string(where) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
    return static_cast<int>(e);
}
int bad_arg(const char* where, const char* what) {
    grid_capi_last_error() = std::string(where) + ": " + what;
    return static_cast<int>(cudaErrorInvalidValue);
}
#define GRID_CHECK_ARGS(FN, COND, NEEDS_SCRATCH)                                        \
    const char* fn = FN;                                                                \
    if (COND) return bad_arg(fn, "bad pointer / stride / num_knots");                   \
    cudaStream_t st = static_cast<cudaStream_t>(s);                                     \
    auto* m = models<T>().get(st, NEEDS_SCRATCH);                                       \
    if (!m) return bad_arg(fn, "no CUDA device / robot model")

inline HostWork* host_work() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) return nullptr;
    return &grid_capi_host_work[dev];
}

// ---- host-buffer entry points.  Two transports, chosen per call:
//  * pipeline (default for pageable or unmapped buffers): chunks of the batch rotate over three streams so that
//    the H2D copy of chunk i+1, the kernel of chunk i and the D2H copy of chunk i-1 overlap (PCIe is full duplex);
//  * zero-copy (GRID_HOST_ZEROCOPY=1, both buffers pinned and mapped into the device address space, which is what
//    cudaHostAlloc / torch pin_memory() give under UVA): ONE kernel launch whose TMA tile loads read the caller's
//    host batch and whose bulk stores write the caller's host result directly over PCIe -- no staging buffers,
//    no copy-engine hand-offs; the double-buffered tile loop overlaps the two PCIe directions by itself.
inline int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}
inline void* mapped_device_pointer(const void* h) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, h) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return a.type == cudaMemoryTypeHost ? a.devicePointer : nullptr;
}
// launch2(d_in, d_in2, d_out, knots, stream); h_in2 (in2_per_knot values per knot, dense) may be null
template <typename T, class Launch2>
int host_pipeline2(const char* fn, const T* h_in, int in_per_knot, const T* h_in2, int in2_per_knot, T* h_out,
                   int out_per_knot, int nk, Launch2 launch2) {
    if (nk < 0 || (nk > 0 && (!h_in || !h_out))) return bad_arg(fn, "bad pointer / num_knots");
    if (nk == 0) return 0;
    HostWork* wp = host_work();
    if (!wp) return bad_arg(fn, "no CUDA device");
    HostWork& w = *wp;
    std::lock_guard<std::mutex> lk(w.mu);
    cudaError_t e;
    static const int zero_copy = env_int("GRID_HOST_ZEROCOPY", 0);
    if (zero_copy) {
        T* m_in = static_cast<T*>(mapped_device_pointer(h_in));
        T* m_in2 = h_in2 ? static_cast<T*>(mapped_device_pointer(h_in2)) : nullptr;
        T* m_out = static_cast<T*>(mapped_device_pointer(h_out));
        if (m_in && m_out && (!h_in2 || m_in2)) {
            if ((e = w.ensure(0, 0, 0)) != cudaSuccess) return fail(e, fn);
            int rc = launch2(m_in, m_in2, m_out, nk, w.streams[0]);
            if (rc) return rc;
            return fail(cudaStreamSynchronize(w.streams[0]), fn);
        }
    }
    e = w.ensure(size_t(nk) * in_per_knot * sizeof(T), h_in2 ? size_t(nk) * in2_per_knot * sizeof(T) : 0,
                 size_t(nk) * out_per_knot * sizeof(T));
    if (e != cudaSuccess) return fail(e, fn);
    // Chunk sizes are multiples of the 32-knot tile (every chunk stays TMA-aligned).  The step is bound by the
    // device->host copies: 25.7 MB per iiwa14 step leave in 0.52 ms as ONE copy with the input copy running beside it and
    // in 0.54 ms as eight (tools/d2h_probe.py, profiles/r02v_d2h_probe.log); the pipeline below takes 0.557 ms with equal
    // chunks of 8192 knots (GRID_HOST_CHUNK).  GRID_HOST_CHUNK_DOUBLING=1: a 2048-knot first chunk, then doubling up to
    // 32 768 -- measured within 2-3 % either way (profiles/r02u_e2e_sweep.log), so equal chunks stay the default.
    static const int fixed_chunk = env_int("GRID_HOST_CHUNK_DOUBLING", 0) ? 0 : ((env_int("GRID_HOST_CHUNK", 8192) + 31) / 32) * 32;
    T* d_in = static_cast<T*>(w.d_in);
    T* d_in2 = h_in2 ? static_cast<T*>(w.d_in2) : nullptr;
    T* d_out = static_cast<T*>(w.d_out);
    int i = 0;
    int chunk = fixed_chunk > 0 ? fixed_chunk : 2048;
    for (int k0 = 0; k0 < nk; k0 += chunk, i++, chunk = fixed_chunk > 0 ? fixed_chunk : (chunk < 32768 ? chunk * 2 : chunk)) {
        const int c = nk - k0 < chunk ? nk - k0 : chunk;
        cudaStream_t st = w.streams[i % 3];
        e = cudaMemcpyAsync(d_in + size_t(k0) * in_per_knot, h_in + size_t(k0) * in_per_knot,
                            size_t(c) * in_per_knot * sizeof(T), cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) return fail(e, fn);
        if (h_in2) {
            e = cudaMemcpyAsync(d_in2 + size_t(k0) * in2_per_knot, h_in2 + size_t(k0) * in2_per_knot,
                                size_t(c) * in2_per_knot * sizeof(T), cudaMemcpyHostToDevice, st);
            if (e != cudaSuccess) return fail(e, fn);
        }
        int rc = launch2(d_in + size_t(k0) * in_per_knot, h_in2 ? d_in2 + size_t(k0) * in2_per_knot : nullptr,
                         d_out + size_t(k0) * out_per_knot, c, st);
        if (rc) return rc;
        e = cudaMemcpyAsync(h_out + size_t(k0) * out_per_knot, d_out + size_t(k0) * out_per_knot,
                            size_t(c) * out_per_knot * sizeof(T), cudaMemcpyDeviceToHost, st);
        if (e != cudaSuccess) return fail(e, fn);
    }
    for (int s = 0; s < 3; s++)
        if ((e = cudaStreamSynchronize(w.streams[s])) != cudaSuccess) return fail(e, fn);
    return 0;
}
template <typename T, class Launch>
int host_pipeline(const char* fn, const T* h_in, int in_per_knot, T* h_out, int out_per_knot, int nk, Launch launch) {
    return host_pipeline2<T>(fn, h_in, in_per_knot, static_cast<const T*>(nullptr), 0, h_out, out_per_knot, nk,
        [&launch](const T* di, const T*, T* dout, int c, cudaStream_t st) { return launch(di, dout, c, st); });
}

#define GRID_INFO(CFG, ALGO, KPTR)                                                                                \
    {                                                                                                             \
        if (what == 0) return long(grid::CFG<T>::SMEM_BYTES);                                                     \
        if (what == 1) return long(grid::CFG<T>::NT);                                                             \
        if (what == 2) return long(grid::ALGO<T>::SLOTS);                                                         \
        if (what == 3) return long(grid::CFG<T>::STATE_SMEM);                                                     \
        if (what == 4) return long(grid::CFG<T>::STAGED);                                                         \
        if (what == 5) return long(grid_rt::persistent_blocks(reinterpret_cast<const void*>(KPTR),                \
                                                              grid::CFG<T>::NT, grid::CFG<T>::SMEM_BYTES, 1 << 20)); \
        return -1;                                                                                                \
    }

}  // namespace

// ---- per-part kernel introspection, dispatched from grid_kernel_info (defined in the core f32 part)
#define DECL_INFO(NAME) HIDDEN long grid_capi_info_##NAME##_f32(int algo, int what); \
                        HIDDEN long grid_capi_info_##NAME##_f64(int algo, int what);
DECL_INFO(core) DECL_INFO(id_du) DECL_INFO(fd_du) DECL_INFO(fd_du_qm) DECL_INFO(aba_crba) DECL_INFO(idsva) DECL_INFO(fdsva) DECL_INFO(ee) DECL_INFO(fd_lin)

// Everything below is instantiated for the dtypes of this translation unit through GRID_FOR_DTYPES(X):
#if (GRID_DTYPES & 1) && (GRID_DTYPES & 2)
#define GRID_FOR_DTYPES(X) X(f32, float) X(f64, double)
#elif (GRID_DTYPES & 1)
#define GRID_FOR_DTYPES(X) X(f32, float)
#else
#define GRID_FOR_DTYPES(X) X(f64, double)
#endif

// =============================================================== core: ID, MINV, FD
#if HAS(PART_CORE)
namespace {
template <typename T>
int inverse_dynamics_impl(const T* d_q_qd, int stride, const T* d_qdd, T* d_c, int nk, T g, void* s) {
    GRID_CHECK_ARGS("grid_inverse_dynamics", nk < 0 || stride < 2 * N || (nk > 0 && (!d_q_qd || !d_c)),
                    grid::inverse_dynamics_cfg<T>::NEEDS_SCRATCH || grid::inverse_dynamics_qdd_cfg<T>::NEEDS_SCRATCH);
    return fail(d_qdd ? grid::inverse_dynamics_qdd_launch<T>(d_c, d_q_qd, stride, d_qdd, m, g, nk, st)
                      : grid::inverse_dynamics_launch<T>(d_c, d_q_qd, stride, m, g, nk, st), fn);
}
template <typename T>
int inverse_dynamics_vaf_impl(const T* d_q_qd, int stride, const T* d_qdd, T* d_vaf, int nk, T g, void* s) {
    GRID_CHECK_ARGS("grid_inverse_dynamics_vaf", nk < 0 || stride < 2 * N || (nk > 0 && (!d_q_qd || !d_vaf)),
                    grid::inverse_dynamics_vaf_cfg<T>::NEEDS_SCRATCH || grid::inverse_dynamics_vaf_qdd_cfg<T>::NEEDS_SCRATCH);
    return fail(d_qdd ? grid::inverse_dynamics_vaf_qdd_launch<T>(d_vaf, d_q_qd, stride, d_qdd, m, g, nk, st)
                      : grid::inverse_dynamics_vaf_launch<T>(d_vaf, d_q_qd, stride, m, g, nk, st), fn);
}
template <typename T>
int direct_minv_impl(const T* d_q, int stride, T* d_Minv, int nk, void* s) {
    GRID_CHECK_ARGS("grid_direct_minv", nk < 0 || stride < N || (nk > 0 && (!d_q || !d_Minv)), grid::direct_minv_cfg<T>::NEEDS_SCRATCH);
    return fail(grid::direct_minv_launch<T>(d_Minv, d_q, stride, m, nk, st), fn);
}
template <typename T>
int forward_dynamics_impl(const T* d_in, int stride, T* d_qdd, int nk, T g, void* s) {
    GRID_CHECK_ARGS("grid_forward_dynamics", nk < 0 || stride < 3 * N || (nk > 0 && (!d_in || !d_qdd)), grid::forward_dynamics_cfg<T>::NEEDS_SCRATCH);
    return fail(grid::forward_dynamics_launch<T>(d_qdd, d_in, stride, m, g, nk, st), fn);
}
template <typename T>
long info_core(int algo, int what) {
    using namespace grid;
    using RM = const robotModel<T>*;
    switch (algo) {
        case 0: GRID_INFO(inverse_dynamics_cfg, inverse_dynamics_algo,
                          static_cast<void (*)(T*, const T*, const int, RM, const T, const int)>(inverse_dynamics_kernel<T>))
        case 1: GRID_INFO(inverse_dynamics_qdd_cfg, inverse_dynamics_qdd_algo,
                          static_cast<void (*)(T*, const T*, const int, const T*, RM, const T, const int)>(inverse_dynamics_kernel<T>))
        case 2: GRID_INFO(direct_minv_cfg, direct_minv_algo,
                          static_cast<void (*)(T*, const T*, const int, RM, const int)>(direct_minv_kernel<T>))
        case 3: GRID_INFO(forward_dynamics_cfg, forward_dynamics_algo,
                          static_cast<void (*)(T*, const T*, const int, RM, const T, const int)>(forward_dynamics_kernel<T>))
    }
    return -1;
}
}  // namespace
#define X_CORE(SUF, T)                                                                                             \
    long grid_capi_info_core_##SUF(int algo, int what) { return info_core<T>(algo, what); }                        \
    extern "C" int grid_inverse_dynamics_##SUF(const T* a, int st, const T* b, T* o, int nk, T g, void* s) {      \
        return inverse_dynamics_impl<T>(a, st, b, o, nk, g, s); }                                                  \
    extern "C" int grid_inverse_dynamics_vaf_##SUF(const T* a, int st, const T* b, T* o, int nk, T g, void* s) {  \
        return inverse_dynamics_vaf_impl<T>(a, st, b, o, nk, g, s); }                                              \
    extern "C" int grid_direct_minv_##SUF(const T* a, int st, T* o, int nk, void* s) {                             \
        return direct_minv_impl<T>(a, st, o, nk, s); }                                                             \
    extern "C" int grid_forward_dynamics_##SUF(const T* a, int st, T* o, int nk, T g, void* s) {                   \
        return forward_dynamics_impl<T>(a, st, o, nk, g, s); }                                                     \
    extern "C" int grid_direct_minv_host_##SUF(const T* h_q, T* h_Minv, int nk) {                                  \
        return host_pipeline<T>("grid_direct_minv_host", h_q, N, h_Minv, N * N, nk,                                \
            [](const T* di, T* dout, int c, cudaStream_t st) { return direct_minv_impl<T>(di, N, dout, c, st); }); }
GRID_FOR_DTYPES(X_CORE)

#if (GRID_DTYPES & 1)
extern "C" {
int grid_num_joints(void) { return N; }
int grid_num_ees(void) { return grid::NUM_EES; }
const char* grid_robot_name(void) { return GRID_ROBOT_NAME; }
const char* grid_last_error(void) { return grid_capi_last_error().c_str(); }
int grid_abi_version(void) { return 7; }
int grid_has_second_order(void) { return GRID_HAS_SECOND_ORDER; }
int grid_has_ee_hessian(void) { return GRID_HAS_EE_HESSIAN; }
const char* grid_fixed_target_name(void) { return GRID_FIXED_TARGET_NAME; }
long grid_kernel_info(int algo, int is_f64, int what) {
#define DISPATCH(NAME) return is_f64 ? grid_capi_info_##NAME##_f64(algo, what) : grid_capi_info_##NAME##_f32(algo, what);
    switch (algo) {
        case 0: case 1: case 2: case 3: DISPATCH(core)
        case 4: case 5: DISPATCH(id_du)
        case 6: DISPATCH(fd_du)
        case 7: DISPATCH(fd_du_qm)
        case 8: case 9: DISPATCH(aba_crba)
        case 10: DISPATCH(idsva)
        case 11: DISPATCH(fdsva)
        case 12: case 13: case 14: DISPATCH(ee)
        case 15: DISPATCH(fd_lin)
    }
    return -1;
}
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
    HostWork* wp = host_work();
    if (!wp) return bad_arg(fn, "no CUDA device");
    HostWork& w = *wp;
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
    {
        int cur = 0;
        cudaGetDevice(&cur);
        for (int d = 0; d < MAX_DEVICES; d++)
            if (grid_capi_host_work[d].have_streams || grid_capi_host_work[d].d_in || grid_capi_host_work[d].d_out) {
                std::lock_guard<std::mutex> lk(grid_capi_host_work[d].mu);
                cudaSetDevice(d);
                grid_capi_host_work[d].free_all();
            }
        cudaSetDevice(cur);
    }
    grid_capi_models_f32.release();
    grid_capi_models_f64.release();
    return 0;
}
}  // extern "C"
#endif
#endif  // PART_CORE

// =============================================================== grad ID
#if HAS(PART_ID_DU)
namespace {
template <typename T>
int inverse_dynamics_gradient_impl(const T* d_q_qd, int stride, const T* d_qdd, T* d_out, int nk, T g, void* s, int form = -1) {
    GRID_CHECK_ARGS("grid_inverse_dynamics_gradient", nk < 0 || stride < 2 * N || (nk > 0 && (!d_q_qd || !d_out)),
                    grid::inverse_dynamics_gradient_cfg<T>::NEEDS_SCRATCH || grid::inverse_dynamics_gradient_qdd_cfg<T>::NEEDS_SCRATCH);
    return fail(d_qdd ? grid::inverse_dynamics_gradient_qdd_launch<T>(d_out, d_q_qd, stride, d_qdd, m, g, nk, st, form)
                      : grid::inverse_dynamics_gradient_launch<T>(d_out, d_q_qd, stride, m, g, nk, st, form), fn);
}
template <typename T>
long info_id_du(int algo, int what) {
    using namespace grid;
    using RM = const robotModel<T>*;
    switch (algo) {
        case 4: GRID_INFO(inverse_dynamics_gradient_cfg, inverse_dynamics_gradient_algo,
                          static_cast<void (*)(T*, const T*, const int, RM, const T, const int)>(inverse_dynamics_gradient_kernel<T>))
        case 5: GRID_INFO(inverse_dynamics_gradient_qdd_cfg, inverse_dynamics_gradient_qdd_algo,
                          static_cast<void (*)(T*, const T*, const int, const T*, RM, const T, const int)>(inverse_dynamics_gradient_kernel<T>))
    }
    return -1;
}
}  // namespace
#define X_ID_DU(SUF, T)                                                                                            \
    long grid_capi_info_id_du_##SUF(int algo, int what) { return info_id_du<T>(algo, what); }                      \
    extern "C" int grid_inverse_dynamics_gradient_##SUF(const T* a, int st, const T* b, T* o, int nk, T g, void* s) { \
        return inverse_dynamics_gradient_impl<T>(a, st, b, o, nk, g, s); }                                         \
    extern "C" int grid_inverse_dynamics_gradient_host_##SUF(const T* h_q_qd, const T* h_qdd, T* h_out, int nk, T g) { \
        /* one kernel form for the whole batch (as the device entry point picks it), passed to every chunk */      \
        const int form = h_qdd ? grid::inverse_dynamics_gradient_qdd_form_for<T>(nk)                               \
                               : grid::inverse_dynamics_gradient_form_for<T>(nk);                                  \
        return host_pipeline2<T>("grid_inverse_dynamics_gradient_host", h_q_qd, 2 * N, h_qdd, N, h_out, 2 * N * N, nk, \
            [g, form](const T* di, const T* dq, T* dout, int c, cudaStream_t st) {                                 \
                return inverse_dynamics_gradient_impl<T>(di, 2 * N, dq, dout, c, g, st, form); }); }
GRID_FOR_DTYPES(X_ID_DU)
#endif

// =============================================================== grad FD (the metric kernel)
#if HAS(PART_FD_DU)
namespace {
template <typename T>
int forward_dynamics_gradient_impl(const T* d_in, int stride, T* d_out, int nk, T g, void* s, int form = -1) {
    GRID_CHECK_ARGS("grid_forward_dynamics_gradient", nk < 0 || stride < 3 * N || (nk > 0 && (!d_in || !d_out)),
                    grid::forward_dynamics_gradient_cfg<T>::NEEDS_SCRATCH);
    return fail(grid::forward_dynamics_gradient_launch<T>(d_out, d_in, stride, m, g, nk, st, form), fn);
}
template <typename T>
long info_fd_du(int algo, int what) {
    using namespace grid;
    using RM = const robotModel<T>*;
    if (algo == 6) GRID_INFO(forward_dynamics_gradient_cfg, forward_dynamics_gradient_algo,
                             static_cast<void (*)(T*, const T*, const int, RM, const T, const int)>(forward_dynamics_gradient_kernel<T>))
    return -1;
}
}  // namespace
#define X_FD_DU(SUF, T)                                                                                            \
    long grid_capi_info_fd_du_##SUF(int algo, int what) { return info_fd_du<T>(algo, what); }                      \
    extern "C" int grid_forward_dynamics_gradient_##SUF(const T* a, int st, T* o, int nk, T g, void* s) {          \
        return forward_dynamics_gradient_impl<T>(a, st, o, nk, g, s); }                                            \
    extern "C" int grid_forward_dynamics_gradient_host_##SUF(const T* h_in, T* h_out, int nk, T g) {               \
        /* one kernel form for the whole batch, so that the chunked result equals the device entry point's */     \
        const int form = grid::forward_dynamics_gradient_form_for<T>(nk);                                          \
        return host_pipeline<T>("grid_forward_dynamics_gradient_host", h_in, 3 * N, h_out, 2 * N * N, nk,          \
            [g, form](const T* di, T* dout, int c, cudaStream_t st) {                                              \
                return forward_dynamics_gradient_impl<T>(di, 3 * N, dout, c, g, st, form); }); }
GRID_FOR_DTYPES(X_FD_DU)
#endif

// =============================================================== grad FD + Minv + qdd
#if HAS(PART_FD_LIN)
namespace {
template <typename T>
int forward_dynamics_linearization_impl(const T* d_in, int stride, T* d_out, int nk, T g, void* s, int form = -1) {
    GRID_CHECK_ARGS("grid_forward_dynamics_linearization", nk < 0 || stride < 3 * N || (nk > 0 && (!d_in || !d_out)),
                    grid::forward_dynamics_linearization_cfg<T>::NEEDS_SCRATCH);
    return fail(grid::forward_dynamics_linearization_launch<T>(d_out, d_in, stride, m, g, nk, st, form), fn);
}
template <typename T>
long info_fd_lin(int algo, int what) {
    using namespace grid;
    using RM = const robotModel<T>*;
    if (algo == 15) GRID_INFO(forward_dynamics_linearization_cfg, forward_dynamics_linearization_algo,
                              static_cast<void (*)(T*, const T*, const int, RM, const T, const int)>(forward_dynamics_linearization_kernel<T>))
    return -1;
}
}  // namespace
#define X_FD_LIN(SUF, T)                                                                                           \
    long grid_capi_info_fd_lin_##SUF(int algo, int what) { return info_fd_lin<T>(algo, what); }                    \
    extern "C" int grid_forward_dynamics_linearization_##SUF(const T* a, int st, T* o, int nk, T g, void* s) {     \
        return forward_dynamics_linearization_impl<T>(a, st, o, nk, g, s); }                                       \
    extern "C" int grid_forward_dynamics_linearization_host_##SUF(const T* h_in, T* h_out, int nk, T g) {          \
        const int form = grid::forward_dynamics_linearization_form_for<T>(nk);                                     \
        return host_pipeline<T>("grid_forward_dynamics_linearization_host", h_in, 3 * N, h_out, 3 * N * N + N, nk, \
            [g, form](const T* di, T* dout, int c, cudaStream_t st) {                                              \
                return forward_dynamics_linearization_impl<T>(di, 3 * N, dout, c, g, st, form); }); }
GRID_FOR_DTYPES(X_FD_LIN)
#endif

#if HAS(PART_FD_DU_QM)
namespace {
template <typename T>
int forward_dynamics_gradient_qdd_minv_impl(const T* d_q_qd, int stride, const T* d_qdd, const T* d_Minv, T* d_out,
                                            int nk, T g, void* s) {
    GRID_CHECK_ARGS("grid_forward_dynamics_gradient_qdd_minv",
                    nk < 0 || stride < 2 * N || (nk > 0 && (!d_q_qd || !d_qdd || !d_Minv || !d_out)),
                    grid::forward_dynamics_gradient_qdd_minv_cfg<T>::NEEDS_SCRATCH);
    return fail(grid::forward_dynamics_gradient_qdd_minv_launch<T>(d_out, d_q_qd, stride, d_qdd, d_Minv, m, g, nk, st), fn);
}
template <typename T>
long info_fd_du_qm(int algo, int what) {
    using namespace grid;
    using RM = const robotModel<T>*;
    if (algo == 7) GRID_INFO(forward_dynamics_gradient_qdd_minv_cfg, forward_dynamics_gradient_qdd_minv_algo,
                             static_cast<void (*)(T*, const T*, const int, const T*, const T*, RM, const T, const int)>(forward_dynamics_gradient_kernel<T>))
    return -1;
}
}  // namespace
#define X_FD_DU_QM(SUF, T)                                                                                         \
    long grid_capi_info_fd_du_qm_##SUF(int algo, int what) { return info_fd_du_qm<T>(algo, what); }                \
    extern "C" int grid_forward_dynamics_gradient_qdd_minv_##SUF(const T* a, int st, const T* b, const T* c, T* o, \
                                                                 int nk, T g, void* s) {                           \
        return forward_dynamics_gradient_qdd_minv_impl<T>(a, st, b, c, o, nk, g, s); }
GRID_FOR_DTYPES(X_FD_DU_QM)
#endif

// =============================================================== ABA, CRBA
#if HAS(PART_ABA_CRBA)
namespace {
template <typename T>
int aba_impl(const T* d_in, int stride, T* d_qdd, int nk, T g, void* s) {
    GRID_CHECK_ARGS("grid_aba", nk < 0 || stride < 3 * N || (nk > 0 && (!d_in || !d_qdd)), grid::aba_cfg<T>::NEEDS_SCRATCH);
    return fail(grid::aba_launch<T>(d_qdd, d_in, stride, m, g, nk, st), fn);
}
template <typename T>
int crba_impl(const T* d_q, int stride, T* d_M, int nk, void* s) {
    GRID_CHECK_ARGS("grid_crba", nk < 0 || stride < N || (nk > 0 && (!d_q || !d_M)), grid::crba_cfg<T>::NEEDS_SCRATCH);
    return fail(grid::crba_launch<T>(d_M, d_q, stride, m, T(0), nk, st), fn);
}
template <typename T>
long info_aba_crba(int algo, int what) {
    using namespace grid;
    using RM = const robotModel<T>*;
    switch (algo) {
        case 8: GRID_INFO(aba_cfg, aba_algo, static_cast<void (*)(T*, const T*, const int, RM, const T, const int)>(aba_kernel<T>))
        case 9: GRID_INFO(crba_cfg, crba_algo, static_cast<void (*)(T*, const T*, const int, RM, const T, const int)>(crba_kernel<T>))
    }
    return -1;
}
}  // namespace
#define X_ABA_CRBA(SUF, T)                                                                                         \
    long grid_capi_info_aba_crba_##SUF(int algo, int what) { return info_aba_crba<T>(algo, what); }                \
    extern "C" int grid_aba_##SUF(const T* a, int st, T* o, int nk, T g, void* s) { return aba_impl<T>(a, st, o, nk, g, s); } \
    extern "C" int grid_crba_##SUF(const T* a, int st, T* o, int nk, void* s) { return crba_impl<T>(a, st, o, nk, s); } \
    extern "C" int grid_aba_host_##SUF(const T* h_in, T* h_qdd, int nk, T g) {                                     \
        return host_pipeline<T>("grid_aba_host", h_in, 3 * N, h_qdd, N, nk,                                        \
            [g](const T* di, T* dout, int c, cudaStream_t st) { return aba_impl<T>(di, 3 * N, dout, c, g, st); }); } \
    extern "C" int grid_crba_host_##SUF(const T* h_q, T* h_M, int nk) {                                            \
        return host_pipeline<T>("grid_crba_host", h_q, N, h_M, N * N, nk,                                          \
            [](const T* di, T* dout, int c, cudaStream_t st) { return crba_impl<T>(di, N, dout, c, st); }); }
GRID_FOR_DTYPES(X_ABA_CRBA)
#endif

// =============================================================== second order (robots up to 8 joints)
#define X_SO(NAME, LAUNCH, CFG, ALGO, KERNEL, IDX)                                                                 \
    namespace {                                                                                                    \
    template <typename T> int NAME##_impl(const T* d_in, int stride, T* d_out, int nk, T g, void* s) {            \
        GRID_CHECK_ARGS("grid_" #NAME, nk < 0 || stride < 3 * N || (nk > 0 && (!d_in || !d_out)),                  \
                        grid::CFG<T>::NEEDS_SCRATCH);                                                              \
        return fail(grid::LAUNCH<T>(d_out, d_in, stride, m, g, nk, st), fn);                                       \
    }                                                                                                              \
    template <typename T> long info_##NAME(int algo, int what) {                                                   \
        using namespace grid;                                                                                      \
        using RM = const robotModel<T>*;                                                                           \
        if (algo == IDX) GRID_INFO(CFG, ALGO, static_cast<void (*)(T*, const T*, const int, RM, const T, const int)>(KERNEL<T>)) \
        return -1;                                                                                                 \
    }                                                                                                              \
    }
#define X_SO_STUB(NAME, SUF, T)                                                                                    \
    long grid_capi_info_##NAME##_##SUF(int, int) { return -1; }                                                    \
    extern "C" int grid_##NAME##_so_host_##SUF(const T*, T*, int, T) {                                             \
        grid_capi_last_error() = "grid_" #NAME "_so_host: this library was built without second-order kernels";    \
        return static_cast<int>(cudaErrorNotSupported); }                                                          \
    extern "C" int grid_##NAME##_so_##SUF(const T*, int, T*, int, T, void*) {                                      \
        grid_capi_last_error() = "grid_" #NAME "_so: this library was built without second-order kernels";         \
        return static_cast<int>(cudaErrorNotSupported); }

#if HAS(PART_IDSVA)
#if GRID_HAS_SECOND_ORDER
X_SO(idsva_so, idsva_so_launch, idsva_so_cfg, idsva_so_algo, idsva_so_kernel, 10)
#define X_IDSVA(SUF, T)                                                                                            \
    long grid_capi_info_idsva_##SUF(int algo, int what) { return info_idsva_so<T>(algo, what); }                   \
    extern "C" int grid_idsva_so_##SUF(const T* a, int st, T* o, int nk, T g, void* s) { return idsva_so_impl<T>(a, st, o, nk, g, s); } \
    extern "C" int grid_idsva_so_host_##SUF(const T* h_in, T* h_out, int nk, T g) {                                \
        return host_pipeline<T>("grid_idsva_so_host", h_in, 3 * N, h_out, 4 * N * N * N, nk,                       \
            [g](const T* di, T* dout, int c, cudaStream_t st) { return idsva_so_impl<T>(di, 3 * N, dout, c, g, st); }); }
#else
#define X_IDSVA(SUF, T) X_SO_STUB(idsva, SUF, T)
#endif
GRID_FOR_DTYPES(X_IDSVA)
#endif

#if HAS(PART_FDSVA)
#if GRID_HAS_SECOND_ORDER
X_SO(fdsva_so, fdsva_so_launch, fdsva_so_cfg, fdsva_so_algo, fdsva_so_kernel, 11)
#define X_FDSVA(SUF, T)                                                                                            \
    long grid_capi_info_fdsva_##SUF(int algo, int what) { return info_fdsva_so<T>(algo, what); }                   \
    extern "C" int grid_fdsva_so_##SUF(const T* a, int st, T* o, int nk, T g, void* s) { return fdsva_so_impl<T>(a, st, o, nk, g, s); } \
    extern "C" int grid_fdsva_so_host_##SUF(const T* h_in, T* h_out, int nk, T g) {                                \
        return host_pipeline<T>("grid_fdsva_so_host", h_in, 3 * N, h_out, 4 * N * N * N, nk,                       \
            [g](const T* di, T* dout, int c, cudaStream_t st) { return fdsva_so_impl<T>(di, 3 * N, dout, c, g, st); }); }
#else
#define X_FDSVA(SUF, T) X_SO_STUB(fdsva, SUF, T)
#endif
GRID_FOR_DTYPES(X_FDSVA)
#endif

// =============================================================== end-effector pose, gradient, Hessian
#if HAS(PART_EE)
namespace {
template <typename T>
int ee_pose_impl(const T* d_q, int stride, T* d_out, int nk, void* s) {
    GRID_CHECK_ARGS("grid_end_effector_pose", nk < 0 || stride < N || (nk > 0 && (!d_q || !d_out)), grid::end_effector_pose_cfg<T>::NEEDS_SCRATCH);
    return fail(grid::end_effector_pose_launch<T>(d_out, d_q, stride, m, nk, st), fn);
}
template <typename T>
int ee_grad_impl(const T* d_q, int stride, T* d_out, int nk, void* s) {
    GRID_CHECK_ARGS("grid_end_effector_pose_gradient", nk < 0 || stride < N || (nk > 0 && (!d_q || !d_out)),
                    grid::end_effector_pose_gradient_cfg<T>::NEEDS_SCRATCH);
    return fail(grid::end_effector_pose_gradient_launch<T>(d_out, d_q, stride, m, nk, st), fn);
}
template <typename T>
int ee_hess_impl(const T* d_q, int stride, T* d_grad, T* d_hess, int nk, void* s) {
#if GRID_HAS_EE_HESSIAN
    GRID_CHECK_ARGS("grid_end_effector_pose_gradient_hessian", nk < 0 || stride < N || (nk > 0 && (!d_q || !d_hess)),
                    grid::end_effector_pose_gradient_cfg<T>::NEEDS_SCRATCH || grid::end_effector_pose_gradient_hessian_cfg<T>::NEEDS_SCRATCH);
    if (d_grad) {
        cudaError_t e = grid::end_effector_pose_gradient_launch<T>(d_grad, d_q, stride, m, nk, st);
        if (e != cudaSuccess) return fail(e, fn);
    }
    return fail(grid::end_effector_pose_gradient_hessian_launch<T>(d_hess, d_q, stride, m, nk, st), fn);
#else
    grid_capi_last_error() = "grid_end_effector_pose_gradient_hessian: this library was built without the EE Hessian kernel";
    return static_cast<int>(cudaErrorNotSupported);
#endif
}
template <typename T>
int joint_transforms_impl(const T* d_q, int stride, T* d_out, int nk, void* s) {
    GRID_CHECK_ARGS("grid_joint_transforms", nk < 0 || stride < N || (nk > 0 && (!d_q || !d_out)), grid::joint_transforms_cfg<T>::NEEDS_SCRATCH);
    return fail(grid::joint_transforms_launch<T>(d_out, d_q, stride, m, nk, st), fn);
}
#if GRID_HAS_FIXED_TARGET
template <typename T>
int fixed_pose_impl(const T* d_q, int stride, T* d_out, int nk, void* s) {
    GRID_CHECK_ARGS("grid_fixed_target_pose", nk < 0 || stride < N || (nk > 0 && (!d_q || !d_out)), grid::GRID_FIXED_TARGET_POSE_CFG<T>::NEEDS_SCRATCH);
    return fail(grid::GRID_FIXED_TARGET_POSE_LAUNCH<T>(d_out, d_q, stride, m, nk, st), fn);
}
template <typename T>
int fixed_grad_impl(const T* d_q, int stride, T* d_out, int nk, void* s) {
    GRID_CHECK_ARGS("grid_fixed_target_pose_gradient", nk < 0 || stride < N || (nk > 0 && (!d_q || !d_out)), grid::GRID_FIXED_TARGET_GRADIENT_CFG<T>::NEEDS_SCRATCH);
    return fail(grid::GRID_FIXED_TARGET_GRADIENT_LAUNCH<T>(d_out, d_q, stride, m, nk, st), fn);
}
template <typename T>
int fixed_hess_impl(const T* d_q, int stride, T* d_grad, T* d_hess, int nk, void* s) {
    GRID_CHECK_ARGS("grid_fixed_target_pose_gradient_hessian", nk < 0 || stride < N || (nk > 0 && (!d_q || !d_hess)),
                    grid::GRID_FIXED_TARGET_GRADIENT_CFG<T>::NEEDS_SCRATCH || grid::GRID_FIXED_TARGET_HESSIAN_CFG<T>::NEEDS_SCRATCH);
    if (d_grad) {
        cudaError_t e = grid::GRID_FIXED_TARGET_GRADIENT_LAUNCH<T>(d_grad, d_q, stride, m, nk, st);
        if (e != cudaSuccess) return fail(e, fn);
    }
    return fail(grid::GRID_FIXED_TARGET_HESSIAN_LAUNCH<T>(d_hess, d_q, stride, m, nk, st), fn);
}
#else
inline int no_fixed_target(const char* fn) {
    grid_capi_last_error() = std::string(fn) + ": this library was built without a fixed kinematic target";
    return static_cast<int>(cudaErrorNotSupported);
}
template <typename T> int fixed_pose_impl(const T*, int, T*, int, void*) { return no_fixed_target("grid_fixed_target_pose"); }
template <typename T> int fixed_grad_impl(const T*, int, T*, int, void*) { return no_fixed_target("grid_fixed_target_pose_gradient"); }
template <typename T> int fixed_hess_impl(const T*, int, T*, T*, int, void*) { return no_fixed_target("grid_fixed_target_pose_gradient_hessian"); }
#endif
template <typename T>
long info_ee(int algo, int what) {
    using namespace grid;
    using RM = const robotModel<T>*;
    typedef void (*K)(T*, const T*, const int, RM, const int);
    switch (algo) {
        case 12: GRID_INFO(end_effector_pose_cfg, end_effector_pose_algo, static_cast<K>(end_effector_pose_kernel<T>))
        case 13: GRID_INFO(end_effector_pose_gradient_cfg, end_effector_pose_gradient_algo, static_cast<K>(end_effector_pose_gradient_kernel<T>))
#if GRID_HAS_EE_HESSIAN
        case 14: GRID_INFO(end_effector_pose_gradient_hessian_cfg, end_effector_pose_gradient_hessian_algo, static_cast<K>(end_effector_pose_gradient_hessian_kernel<T>))
#endif
    }
    return -1;
}
}  // namespace
#define X_EE(SUF, T)                                                                                               \
    long grid_capi_info_ee_##SUF(int algo, int what) { return info_ee<T>(algo, what); }                            \
    extern "C" int grid_end_effector_pose_##SUF(const T* a, int st, T* o, int nk, void* s) { return ee_pose_impl<T>(a, st, o, nk, s); } \
    extern "C" int grid_end_effector_pose_gradient_##SUF(const T* a, int st, T* o, int nk, void* s) { return ee_grad_impl<T>(a, st, o, nk, s); } \
    extern "C" int grid_end_effector_pose_gradient_hessian_##SUF(const T* a, int st, T* g, T* h, int nk, void* s) { \
        return ee_hess_impl<T>(a, st, g, h, nk, s); }                                                              \
    extern "C" int grid_joint_transforms_##SUF(const T* a, int st, T* o, int nk, void* s) { return joint_transforms_impl<T>(a, st, o, nk, s); } \
    extern "C" int grid_fixed_target_pose_##SUF(const T* a, int st, T* o, int nk, void* s) { return fixed_pose_impl<T>(a, st, o, nk, s); } \
    extern "C" int grid_fixed_target_pose_gradient_##SUF(const T* a, int st, T* o, int nk, void* s) { return fixed_grad_impl<T>(a, st, o, nk, s); } \
    extern "C" int grid_fixed_target_pose_gradient_hessian_##SUF(const T* a, int st, T* g, T* h, int nk, void* s) { \
        return fixed_hess_impl<T>(a, st, g, h, nk, s); }
GRID_FOR_DTYPES(X_EE)
#endif
