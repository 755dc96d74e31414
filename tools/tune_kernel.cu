// Stand-alone timing of ONE emitted kernel with a given (NT, STATE_SMEM, STAGED) configuration; built by
// tools/tune_config.py against a copy of grid.cuh whose <algo>_cfg struct was rewritten.
//   nvcc -DALGO=forward_dynamics_gradient -DINCOLS=3 -DGRAV=1 -I<dir with grid.cuh> tools/tune_kernel.cu
#include "grid.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>
#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)
#define STR_(x) #x
#define STR(x) STR_(x)
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: CUDA error %s\n", TAG, cudaGetErrorString(e_)); return 1; } } while (0)

int main(int argc, char** argv) {
    using namespace grid;
    const int B = argc > 1 ? atoi(argv[1]) : 8192;
    const int steps = argc > 2 ? atoi(argv[2]) : 20;
    const int n = NUM_JOINTS, in_w = INCOLS * n, out_w = CAT(ALGO, _algo)<float>::OUT;
    const int nsets = 3;
    robotModel<float>* rm = init_robotModel<float>();
    std::vector<float> h(size_t(B) * in_w);
    srand(1);
    for (auto& v : h) v = float(rand()) / RAND_MAX * 2.f - 1.f;
    float *in[nsets], *out[nsets];
    for (int s = 0; s < nsets; s++) {
        CK(cudaMalloc(&in[s], h.size() * 4));
        CK(cudaMalloc(&out[s], size_t(B) * out_w * 4));
        CK(cudaMemcpy(in[s], h.data(), h.size() * 4, cudaMemcpyHostToDevice));
    }
    cudaStream_t st;
    CK(cudaStreamCreate(&st));
    auto launch = [&](int s) {
#if GRAV
        return CAT(ALGO, _launch)<float>(out[s], in[s], in_w, rm, 9.81f, B, st);
#else
        return CAT(ALGO, _launch)<float>(out[s], in[s], in_w, rm, B, st);
#endif
    };
    for (int i = 0; i < 3; i++) CK(launch(i % nsets));
    CK(cudaStreamSynchronize(st));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0, st);
        for (int i = 0; i < steps; i++) CK(launch(i % nsets));
        cudaEventRecord(e1, st);
        CK(cudaStreamSynchronize(st));
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms / steps < best) best = ms / steps;
    }
    std::vector<float> ho(size_t(B) * out_w);
    CK(cudaMemcpy(ho.data(), out[0], ho.size() * 4, cudaMemcpyDeviceToHost));
    double cs = 0; for (float v : ho) cs += v < 0 ? -v : v;
    printf("%s B=%d us=%.2f checksum=%.6e\n", TAG, B, best * 1e3, cs);
    return 0;
}
