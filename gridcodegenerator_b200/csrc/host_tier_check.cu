// host_tier_check.cu -- a user program written against the emitted grid.cuh exactly as the reference documents it
// (reference GRiDCodeGenerator.py:312-380 banner; host tier algorithms/_forward_dynamics_gradient.py:186-249,
// algorithms/_inverse_dynamics.py:452-511): init_robotModel / init_grid / init_gridData<T,N> -> fill hd_data->h_* ->
// ALGORITHM<T,FLAGS>(hd_data, d_robotModel, gravity, num_timesteps, block_dimms, thread_dimms, streams) -> read
// hd_data->h_* -> close_grid.  It instantiates and RUNS the host tier (plain, _compute_only, _single_timing) so that the
// header's user surface is type-checked by nvcc and its results can be compared with the C ABI
// (tests/test_gpu_host_tier.py does that bit for bit).
//
//   host_tier_check <input.bin> <output.bin>
// input : KNOTS x 4n floats, rows [q | qd | u | qdd]      output: the sections listed in main(), concatenated
#include "grid.cuh"

#include <cstdio>
#include <cstring>
#include <vector>

#ifndef HOST_TIER_KNOTS
#define HOST_TIER_KNOTS 1000
#endif

using namespace grid;
typedef float T;

static std::vector<T> g_out;
static void section(const T* p, size_t count) { g_out.insert(g_out.end(), p, p + count); }

// ---- user kernels written against the REFERENCE's device tier (banner GRiDCodeGenerator.py:318-352): one thread block
// per knot point (grid-strided), inputs and results in shared memory, X_device<T>(s_out, s_q, s_qd, ..., d_robotModel,
// gravity) called by every thread of the block.  d_x rows are [q | qd | u | qdd].
constexpr int NJ = NUM_JOINTS;
template <int WHICH>
__global__ void user_device_tier(T* d_out, const int out_n, const T* d_x, const T* d_Minv, const robotModel<T>* d_robotModel,
                                 const T gravity, const int N) {
    __shared__ T s_x[4 * NJ];
    __shared__ T s_Minv[NJ * NJ];
    extern __shared__ T s_out[];
    const int tid = threadIdx.x + threadIdx.y * blockDim.x, nthreads = blockDim.x * blockDim.y;
    const T *s_q = s_x, *s_qd = s_x + NJ, *s_u = s_x + 2 * NJ, *s_qdd = s_x + 3 * NJ;
    for (int k = blockIdx.x; k < N; k += gridDim.x) {
        for (int i = tid; i < 4 * NJ; i += nthreads) s_x[i] = d_x[size_t(k) * 4 * NJ + i];
        for (int i = tid; i < NJ * NJ; i += nthreads) s_Minv[i] = d_Minv[size_t(k) * NJ * NJ + i];
        __syncthreads();
        if (WHICH == 0) inverse_dynamics_device<T>(s_out, s_q, s_qd, d_robotModel, gravity);
        if (WHICH == 1) inverse_dynamics_device<T>(s_out, s_q, s_qd, s_qdd, d_robotModel, gravity);
        if (WHICH == 2) direct_minv_device<T>(s_out, s_q, d_robotModel);
        if (WHICH == 3) forward_dynamics_device<T>(s_out, s_q, s_qd, s_u, d_robotModel, gravity);
        if (WHICH == 4) inverse_dynamics_gradient_device<T>(s_out, s_q, s_qd, s_qdd, d_robotModel, gravity);
        if (WHICH == 5) forward_dynamics_gradient_device<T>(s_out, s_q, s_qd, s_u, d_robotModel, gravity);
        if (WHICH == 6) forward_dynamics_gradient_device<T>(s_out, s_q, s_qd, s_qdd, s_Minv, d_robotModel, gravity);
        if (WHICH == 7) end_effector_pose_device<T>(s_out, s_q, d_robotModel);
        if (WHICH == 8) inverse_dynamics_vaf_device<T>(s_out, s_q, s_qd, s_qdd, d_robotModel, gravity);
        __syncthreads();
        for (int i = tid; i < out_n; i += nthreads) d_out[size_t(k) * out_n + i] = s_out[i];
        __syncthreads();
    }
}
template <int WHICH>
static void run_device_tier(int out_n, const T* d_x, const T* d_Minv, const robotModel<T>* d_robotModel, T gravity, int N) {
    T* d_out;
    gpuErrchk(cudaMalloc(&d_out, size_t(N) * out_n * sizeof(T)));
    gpuErrchk(cudaMemset(d_out, 0, size_t(N) * out_n * sizeof(T)));
    user_device_tier<WHICH><<<296, dim3(16, 4), out_n * sizeof(T)>>>(d_out, out_n, d_x, d_Minv, d_robotModel, gravity, N);
    gpuErrchk(cudaGetLastError());
    std::vector<T> h(size_t(N) * out_n);
    gpuErrchk(cudaMemcpy(h.data(), d_out, h.size() * sizeof(T), cudaMemcpyDeviceToHost));
    gpuErrchk(cudaFree(d_out));
    section(h.data(), h.size());
}

int main(int argc, char** argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s input.bin output.bin\n", argv[0]); return 2; }
    constexpr int N = HOST_TIER_KNOTS;
    constexpr int n = NUM_JOINTS;
    std::vector<T> in(size_t(N) * 4 * n);
    FILE* f = fopen(argv[1], "rb");
    if (!f || fread(in.data(), sizeof(T), in.size(), f) != in.size()) { fprintf(stderr, "bad input file\n"); return 2; }
    fclose(f);

    const T gravity = static_cast<T>(9.81);
    const dim3 dimms(1, 1, 1);                       // accepted for source compatibility; the grid is sized by occupancy
    robotModel<T>* d_robotModel = init_robotModel<T>();
    cudaStream_t* streams = init_grid<T>();
    gridData<T>* hd_data = init_gridData<T, N>();

    for (int k = 0; k < N; k++) {
        const T* row = &in[size_t(k) * 4 * n];
        memcpy(&hd_data->h_q_qd_u[size_t(k) * 3 * n], row, 3 * n * sizeof(T));
        memcpy(&hd_data->h_q_qd[size_t(k) * 2 * n], row, 2 * n * sizeof(T));
        memcpy(&hd_data->h_q[size_t(k) * n], row, n * sizeof(T));
        memcpy(&hd_data->h_qdd[size_t(k) * n], row + 3 * n, n * sizeof(T));
    }

    // 0: RNEA without qdd from the [q|qd|u] buffer
    inverse_dynamics<T, false, false>(hd_data, d_robotModel, gravity, N, dimms, dimms, streams);
    section(hd_data->h_c, size_t(N) * n);
    // 1: RNEA with qdd from the compressed [q|qd] buffer
    inverse_dynamics<T, true, true>(hd_data, d_robotModel, gravity, N, dimms, dimms, streams);
    section(hd_data->h_c, size_t(N) * n);
    // 2: grad RNEA with qdd
    inverse_dynamics_gradient<T, true, false>(hd_data, d_robotModel, gravity, N, dimms, dimms, streams);
    section(hd_data->h_dc_du, size_t(N) * 2 * n * n);
    // 3: direct Minv
    direct_minv<T, false>(hd_data, d_robotModel, N, dimms, dimms, streams);
    section(hd_data->h_Minv, size_t(N) * n * n);
    // 4: FD (overwrites h_qdd with qdd = FD(q, qd, u), which USE_QDD_MINV_FLAG consumes below)
    forward_dynamics<T>(hd_data, d_robotModel, gravity, N, dimms, dimms, streams);
    section(hd_data->h_qdd, size_t(N) * n);
    // 5: grad FD, self-contained
    forward_dynamics_gradient<T, false>(hd_data, d_robotModel, gravity, N, dimms, dimms, streams);
    section(hd_data->h_df_du, size_t(N) * 2 * n * n);
    // 6: grad FD with the caller's qdd and Minv (reference USE_QDD_MINV_FLAG)
    forward_dynamics_gradient<T, true>(hd_data, d_robotModel, gravity, N, dimms, dimms, streams);
    section(hd_data->h_df_du, size_t(N) * 2 * n * n);
    // 7: _compute_only: inputs already on the device (left there by 5/6), result fetched by hand
    gpuErrchk(cudaMemset(hd_data->d_df_du, 0, size_t(N) * 2 * n * n * sizeof(T)));
    forward_dynamics_gradient_compute_only<T, false>(hd_data, d_robotModel, gravity, N, dimms, dimms);
    gpuErrchk(cudaMemcpy(hd_data->h_df_du, hd_data->d_df_du, size_t(N) * 2 * n * n * sizeof(T), cudaMemcpyDeviceToHost));
    section(hd_data->h_df_du, size_t(N) * 2 * n * n);
    // 8: _single_timing (prints "Single Call FD_DU ...us")
    forward_dynamics_gradient_single_timing<T, false>(hd_data, d_robotModel, gravity, N, dimms, dimms, streams);
    section(hd_data->h_df_du, size_t(N) * 2 * n * n);
    // 9: grad FD + Minv + qdd in one launch
    forward_dynamics_linearization<T>(hd_data, d_robotModel, gravity, N, dimms, dimms, streams);
    section(hd_data->h_fd_lin, size_t(N) * (3 * n * n + n));
    // 10: ABA, 11: CRBA
    aba<T>(hd_data, d_robotModel, gravity, N, dimms, dimms, streams);
    section(hd_data->h_qdd, size_t(N) * n);
    crba<T, false>(hd_data, d_robotModel, gravity, N, dimms, dimms, streams);
    section(hd_data->h_M, size_t(N) * n * n);
    // 12: end-effector pose, 13: its gradient
    end_effector_pose<T, false>(hd_data, d_robotModel, N, dimms, dimms, streams);
    section(hd_data->h_eePos, size_t(N) * 6 * NUM_EES);
    end_effector_pose_gradient<T, true>(hd_data, d_robotModel, N, dimms, dimms, streams);
    section(hd_data->h_deePos, size_t(N) * 6 * NUM_EES * n);

    // a user kernel launch with the reference's kernel parameter list (b1: kernel tier)
    {
        gpuErrchk(cudaMemset(hd_data->d_c, 0, size_t(N) * n * sizeof(T)));
        auto kern = static_cast<void (*)(T*, const T*, const int, const robotModel<T>*, const T, const int)>(inverse_dynamics_kernel<T>);
        gpuErrchk(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       int(ID_DYNAMIC_SHARED_MEM_COUNT * sizeof(T))));
        kern<<<148, inverse_dynamics_cfg<T>::NT, ID_DYNAMIC_SHARED_MEM_COUNT * sizeof(T)>>>(
            hd_data->d_c, hd_data->d_q_qd_u, 3 * n, d_robotModel, gravity, N);
        gpuErrchk(cudaGetLastError());
        gpuErrchk(cudaMemcpy(hd_data->h_c, hd_data->d_c, size_t(N) * n * sizeof(T), cudaMemcpyDeviceToHost));
        section(hd_data->h_c, size_t(N) * n);    // 14 == 0
    }

    // 15: cumulative joint transforms (the quantity the reference's X_single_thread / X_warp accumulate)
    joint_transforms<T, true>(hd_data, d_robotModel, N, dimms, dimms, streams);
    section(hd_data->h_jointXforms, size_t(N) * 16 * n);

    // 16-24: the reference's DEVICE tier called from block-per-knot user kernels (b1: device tier)
    {
        T* d_x;
        gpuErrchk(cudaMalloc(&d_x, in.size() * sizeof(T)));
        gpuErrchk(cudaMemcpy(d_x, in.data(), in.size() * sizeof(T), cudaMemcpyHostToDevice));
        const T* d_Minv = hd_data->d_Minv;                     // left on the device by section 3
        run_device_tier<0>(n, d_x, d_Minv, d_robotModel, gravity, N);
        run_device_tier<1>(n, d_x, d_Minv, d_robotModel, gravity, N);
        run_device_tier<2>(n * n, d_x, d_Minv, d_robotModel, gravity, N);
        run_device_tier<3>(n, d_x, d_Minv, d_robotModel, gravity, N);
        run_device_tier<4>(2 * n * n, d_x, d_Minv, d_robotModel, gravity, N);
        run_device_tier<5>(2 * n * n, d_x, d_Minv, d_robotModel, gravity, N);
        run_device_tier<6>(2 * n * n, d_x, d_Minv, d_robotModel, gravity, N);
        run_device_tier<7>(6 * NUM_EES, d_x, d_Minv, d_robotModel, gravity, N);
        run_device_tier<8>(18 * n, d_x, d_Minv, d_robotModel, gravity, N);
        gpuErrchk(cudaFree(d_x));
    }

    close_grid<T>(streams, d_robotModel, hd_data);

    f = fopen(argv[2], "wb");
    if (!f || fwrite(g_out.data(), sizeof(T), g_out.size(), f) != g_out.size()) { fprintf(stderr, "cannot write output\n"); return 2; }
    fclose(f);
    printf("host_tier_check: %d knots, %zu values written\n", N, g_out.size());
    return 0;
}
