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

    close_grid<T>(streams, d_robotModel, hd_data);

    f = fopen(argv[2], "wb");
    if (!f || fwrite(g_out.data(), sizeof(T), g_out.size(), f) != g_out.size()) { fprintf(stderr, "cannot write output\n"); return 2; }
    fclose(f);
    printf("host_tier_check: %d knots, %zu values written\n", N, g_out.size());
    return 0;
}
