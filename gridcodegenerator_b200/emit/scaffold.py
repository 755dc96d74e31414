"""File-level scaffolding of grid.cuh: includes, constants, robotModel<T>, gridData<T>, lifecycle.

Mirrors what the reference's driver emits (GRiDCodeGenerator.py:66-306): same constant names,
same struct field names, same lifecycle function names.  Differences (deliberate):
host buffers are pinned (cudaMallocHost) instead of malloc'd, every buffer of gridData is freed by
close_grid, and robotModel<T> carries only the optional global state scratch because the robot's
constants are baked into the emitted code.
"""
from __future__ import annotations

INCLUDES = """#include <algorithm>
#include <assert.h>
#include <stdio.h>
#include <stdlib.h>
#include <time.h>
#include <cuda_runtime.h>
#include <stdint.h>

// single kernel timing helper (reference GRiDCodeGenerator.py:81-82)
#define time_delta_us_timespec(start,end) (1e6*static_cast<double>(end.tv_sec - start.tv_sec)+1e-3*static_cast<double>(end.tv_nsec - start.tv_nsec))
"""

GPU_ERR = """/**
 * Check for runtime errors using CUDA (reference GRiDCodeGenerator.py:273-288).
 * If no error is detected, the function does nothing.
 */
#ifndef gpuErrchk
#define gpuErrchk(ans) { ::gpuAssert((ans), __FILE__, __LINE__); }
__host__ inline
void gpuAssert(cudaError_t code, const char *file, const int line, bool abort=true){
    if (code != cudaSuccess){
        fprintf(stderr,"GPUassert: %s %s %d\\n", cudaGetErrorString(code), file, line);
        if (abort){cudaDeviceReset(); exit(code);}
    }
}
#endif
"""

PRINT_MAT = """template <typename T, int M, int N>
__host__ __device__
void printMat(T *A, int lda){
    for(int i=0; i<M; i++){
        for(int j=0; j<N; j++){printf("%.4f ",A[i + lda*j]);}
        printf("\\n");
    }
}
"""

STRUCTS = """template <typename T>
struct robotModel {
    T *d_scratch;        // global per-thread state scratch, only for algorithms whose state exceeds shared memory
    int scratch_blocks;  // number of CTAs the scratch can serve
    // kept for source compatibility with code written against the reference struct (GRiDCodeGenerator.py:115-119);
    // the robot's constants are immediates in the emitted code, so these are nullptr.
    T *d_XImats;
    int *d_topology_helpers;
};
template <typename T>
struct gridData {
    // GPU INPUTS
    T *d_q_qd_u;
    T *d_q_qd;
    T *d_q;
    // CPU INPUTS (pinned)
    T *h_q_qd_u;
    T *h_q_qd;
    T *h_q;
    // GPU OUTPUTS
    T *d_c;
    T *d_Minv;
    T *d_qdd;
    T *d_M;
    T *d_dc_du;
    T *d_df_du;
    T *d_fd_lin;     // [df_du | Minv | qdd], 3 n^2 + n per knot (forward_dynamics_linearization)
    T *d_eePos;
    T *d_deePos;
    T *d_d2eePos;
    T *d_jointXforms; // 16 n per knot: cumulative 4 x 4 joint transforms (joint_transforms)
    T *d_idsva_so;   // 4 n^3 per knot, allocated only when GRID_HAS_SECOND_ORDER
    T *d_df2;
    // CPU OUTPUTS (pinned)
    T *h_c;
    T *h_Minv;
    T *h_qdd;
    T *h_M;
    T *h_dc_du;
    T *h_df_du;
    T *h_fd_lin;
    T *h_eePos;
    T *h_deePos;
    T *h_d2eePos;
    T *h_jointXforms;
    T *h_idsva_so;
    T *h_df2;
};
"""

LIFECYCLE = """/**
 * Allocates the (device-resident) robot model handle.  The constants of the robot live in the code;
 * the handle owns the optional global state scratch (indexed by CTA: at most ONE scratch-using kernel --
 * <algo>_cfg<T>::NEEDS_SCRATCH -- may be in flight per handle; use one handle per stream for concurrent launches).
 * with_scratch = false gives a handle for kernels that never touch the scratch.
 */
template <typename T>
__host__
robotModel<T>* init_robotModel(bool with_scratch = true) {
    robotModel<T> h_model;
    h_model.d_scratch = nullptr; h_model.scratch_blocks = 0;
    h_model.d_XImats = nullptr; h_model.d_topology_helpers = nullptr;
    const size_t per_block = sizeof(T) == 4 ? GRID_SCRATCH_ELEMS_PER_BLOCK_F32 : GRID_SCRATCH_ELEMS_PER_BLOCK_F64;
    if (per_block > 0 && with_scratch) {
        h_model.scratch_blocks = grid_rt::scratch_capacity(GRID_SCRATCH_BLOCKS_PER_SM);
        if (h_model.scratch_blocks <= 0) gpuErrchk(cudaErrorInvalidDevice);
        gpuErrchk(cudaMalloc((void**)&h_model.d_scratch, per_block*h_model.scratch_blocks*sizeof(T)));
    }
    robotModel<T> *d_robotModel; gpuErrchk(cudaMalloc((void**)&d_robotModel, sizeof(robotModel<T>)));
    gpuErrchk(cudaMemcpy(d_robotModel, &h_model, sizeof(robotModel<T>), cudaMemcpyHostToDevice));
    return d_robotModel;
}

/**
 * Allocates all memory and returns the gridData bundle (reference GRiDCodeGenerator.py:160-213).
 */
template <typename T>
__host__
gridData<T> *init_gridData(const int num_timesteps){
    gridData<T> *hd_data = (gridData<T> *)malloc(sizeof(gridData<T>));
    const size_t N = num_timesteps, n = NUM_JOINTS;
    gpuErrchk(cudaMalloc((void**)&hd_data->d_q_qd_u, 3*n*N*sizeof(T)));
    gpuErrchk(cudaMalloc((void**)&hd_data->d_q_qd, 2*n*N*sizeof(T)));
    gpuErrchk(cudaMalloc((void**)&hd_data->d_q, n*N*sizeof(T)));
    gpuErrchk(cudaMalloc((void**)&hd_data->d_c, n*N*sizeof(T)));
    gpuErrchk(cudaMalloc((void**)&hd_data->d_Minv, n*n*N*sizeof(T)));
    gpuErrchk(cudaMalloc((void**)&hd_data->d_qdd, n*N*sizeof(T)));
    gpuErrchk(cudaMalloc((void**)&hd_data->d_M, n*n*N*sizeof(T)));
    gpuErrchk(cudaMalloc((void**)&hd_data->d_dc_du, 2*n*n*N*sizeof(T)));
    gpuErrchk(cudaMalloc((void**)&hd_data->d_df_du, 2*n*n*N*sizeof(T)));
    gpuErrchk(cudaMalloc((void**)&hd_data->d_fd_lin, (3*n*n + n)*N*sizeof(T)));
    gpuErrchk(cudaMallocHost((void**)&hd_data->h_q_qd_u, 3*n*N*sizeof(T)));
    gpuErrchk(cudaMallocHost((void**)&hd_data->h_q_qd, 2*n*N*sizeof(T)));
    gpuErrchk(cudaMallocHost((void**)&hd_data->h_q, n*N*sizeof(T)));
    gpuErrchk(cudaMallocHost((void**)&hd_data->h_c, n*N*sizeof(T)));
    gpuErrchk(cudaMallocHost((void**)&hd_data->h_Minv, n*n*N*sizeof(T)));
    gpuErrchk(cudaMallocHost((void**)&hd_data->h_qdd, n*N*sizeof(T)));
    gpuErrchk(cudaMallocHost((void**)&hd_data->h_M, n*n*N*sizeof(T)));
    gpuErrchk(cudaMallocHost((void**)&hd_data->h_dc_du, 2*n*n*N*sizeof(T)));
    gpuErrchk(cudaMallocHost((void**)&hd_data->h_df_du, 2*n*n*N*sizeof(T)));
    gpuErrchk(cudaMallocHost((void**)&hd_data->h_fd_lin, (3*n*n + n)*N*sizeof(T)));
    gpuErrchk(cudaMalloc((void**)&hd_data->d_eePos, 6*NUM_EES*N*sizeof(T)));
    gpuErrchk(cudaMalloc((void**)&hd_data->d_deePos, 6*NUM_EES*n*N*sizeof(T)));
    gpuErrchk(cudaMalloc((void**)&hd_data->d_d2eePos, 6*NUM_EES*n*n*N*sizeof(T)));
    gpuErrchk(cudaMallocHost((void**)&hd_data->h_eePos, 6*NUM_EES*N*sizeof(T)));
    gpuErrchk(cudaMallocHost((void**)&hd_data->h_deePos, 6*NUM_EES*n*N*sizeof(T)));
    gpuErrchk(cudaMallocHost((void**)&hd_data->h_d2eePos, 6*NUM_EES*n*n*N*sizeof(T)));
    gpuErrchk(cudaMalloc((void**)&hd_data->d_jointXforms, 16*n*N*sizeof(T)));
    gpuErrchk(cudaMallocHost((void**)&hd_data->h_jointXforms, 16*n*N*sizeof(T)));
    hd_data->d_idsva_so = hd_data->d_df2 = hd_data->h_idsva_so = hd_data->h_df2 = nullptr;
#if GRID_HAS_SECOND_ORDER
    gpuErrchk(cudaMalloc((void**)&hd_data->d_idsva_so, 4*n*n*n*N*sizeof(T)));
    gpuErrchk(cudaMalloc((void**)&hd_data->d_df2, 4*n*n*n*N*sizeof(T)));
    gpuErrchk(cudaMallocHost((void**)&hd_data->h_idsva_so, 4*n*n*n*N*sizeof(T)));
    gpuErrchk(cudaMallocHost((void**)&hd_data->h_df2, 4*n*n*n*N*sizeof(T)));
#endif
    return hd_data;
}
template <typename T, int NUM_TIMESTEPS>
__host__
gridData<T> *init_gridData(){ return init_gridData<T>(NUM_TIMESTEPS); }

/**
 * Allocates the three H2D streams of the reference API (reference GRiDCodeGenerator.py:215-250).
 */
template <typename T>
__host__
cudaStream_t *init_grid(){
    cudaStream_t *streams = (cudaStream_t *)malloc(3*sizeof(cudaStream_t));
    int priority, minPriority, maxPriority;
    gpuErrchk(cudaDeviceGetStreamPriorityRange(&minPriority, &maxPriority));
    for(int i=0; i<3; i++){
        priority = std::min(minPriority+i, maxPriority);
        gpuErrchk(cudaStreamCreateWithPriority(&(streams[i]),cudaStreamNonBlocking,priority));
    }
    return streams;
}

/**
 * Frees the memory used by grid (reference GRiDCodeGenerator.py:252-271).
 */
template <typename T>
__host__
void close_grid(cudaStream_t *streams, robotModel<T> *d_robotModel, gridData<T> *hd_data){
    if (d_robotModel) {
        robotModel<T> h_model;
        gpuErrchk(cudaMemcpy(&h_model, d_robotModel, sizeof(robotModel<T>), cudaMemcpyDeviceToHost));
        if (h_model.d_scratch) { gpuErrchk(cudaFree(h_model.d_scratch)); }
        gpuErrchk(cudaFree(d_robotModel));
    }
    if (hd_data) {
        gpuErrchk(cudaFree(hd_data->d_q_qd_u)); gpuErrchk(cudaFree(hd_data->d_q_qd)); gpuErrchk(cudaFree(hd_data->d_q));
        gpuErrchk(cudaFree(hd_data->d_c)); gpuErrchk(cudaFree(hd_data->d_Minv)); gpuErrchk(cudaFree(hd_data->d_qdd));
        gpuErrchk(cudaFree(hd_data->d_M)); gpuErrchk(cudaFree(hd_data->d_dc_du)); gpuErrchk(cudaFree(hd_data->d_df_du));
        gpuErrchk(cudaFree(hd_data->d_fd_lin)); gpuErrchk(cudaFreeHost(hd_data->h_fd_lin));
        gpuErrchk(cudaFreeHost(hd_data->h_q_qd_u)); gpuErrchk(cudaFreeHost(hd_data->h_q_qd)); gpuErrchk(cudaFreeHost(hd_data->h_q));
        gpuErrchk(cudaFreeHost(hd_data->h_c)); gpuErrchk(cudaFreeHost(hd_data->h_Minv)); gpuErrchk(cudaFreeHost(hd_data->h_qdd));
        gpuErrchk(cudaFreeHost(hd_data->h_M)); gpuErrchk(cudaFreeHost(hd_data->h_dc_du)); gpuErrchk(cudaFreeHost(hd_data->h_df_du));
        gpuErrchk(cudaFree(hd_data->d_eePos)); gpuErrchk(cudaFree(hd_data->d_deePos)); gpuErrchk(cudaFree(hd_data->d_d2eePos));
        gpuErrchk(cudaFreeHost(hd_data->h_eePos)); gpuErrchk(cudaFreeHost(hd_data->h_deePos)); gpuErrchk(cudaFreeHost(hd_data->h_d2eePos));
        gpuErrchk(cudaFree(hd_data->d_jointXforms)); gpuErrchk(cudaFreeHost(hd_data->h_jointXforms));
        if (hd_data->d_idsva_so) { gpuErrchk(cudaFree(hd_data->d_idsva_so)); gpuErrchk(cudaFree(hd_data->d_df2)); }
        if (hd_data->h_idsva_so) { gpuErrchk(cudaFreeHost(hd_data->h_idsva_so)); gpuErrchk(cudaFreeHost(hd_data->h_df2)); }
        free(hd_data);
    }
    if (streams) {
        for(int i=0; i<3; i++){gpuErrchk(cudaStreamDestroy(streams[i]));}
        free(streams);
    }
}
"""
