"""The reference's public code-text helpers (README.md:44-56; helpers/_code_generation_helpers.py:34-190,
helpers/_topology_helpers.py:515-542) as provided by this build (emit/user_helpers.py): a user assembles a
block-per-knot kernel from them on the generator's code buffer; the text must compile with nvcc, and on a GPU it must
do what it says."""
import os
import subprocess
import tempfile

import numpy as np
import pytest

from gridcodegenerator_b200 import GRiDCodeGenerator, load_robot


def _user_kernel_source():
    from gridcodegenerator_b200.emit.user_helpers import mx_device_helpers
    g = GRiDCodeGenerator(load_robot("branched6"))
    g.code_str, g.indent_level = "", 0
    g.gen_add_code_lines(["#include <cstdio>", "#include <cooperative_groups.h>", "namespace cgrps = cooperative_groups;", ""])
    g.code_str += mx_device_helpers()          # in a generated grid.cuh these come with the header
    g.gen_add_func_doc("y[k] = 2 * x[k] on the first half of every row, 3 * x[k] on the second half; z[k] = sum of row k",
                       ["one thread block per knot point (grid-strided), rows of 12 values"],
                       ["d_y is the output", "d_x is the input", "N is the number of rows"], None)
    g.gen_add_code_line("template <typename T>")
    g.gen_add_code_line("__global__ void user_kernel(T *d_y, T *d_z, const T *d_x, const int N) {", True)
    g.gen_add_code_line("__shared__ T s_x[12]; __shared__ T s_y[12]; __shared__ T s_z[1];")
    g.gen_add_parallel_loop("k", "N", block_level=True)
    g.gen_kernel_load_inputs("x", "12", "12")
    g.gen_add_parallel_loop("ind", "12")
    g.gen_add_multi_threaded_select("ind", "<", ["6", "12"], [("T", "scale", ["static_cast<T>(2)", "static_cast<T>(3)"])])
    g.gen_add_code_line("s_y[ind] = scale * s_x[ind];")
    g.gen_add_end_control_flow()
    g.gen_add_sync()
    g.gen_add_serial_ops()
    g.gen_add_code_line("T acc = 0; for (int i = 0; i < 12; i++) { acc += s_x[i]; } s_z[0] = acc;")
    g.gen_add_end_control_flow()
    g.gen_add_sync()
    g.gen_kernel_save_result("y", "12", "12")
    g.gen_kernel_save_result("z", "1", "1")
    g.gen_add_end_control_flow()
    g.gen_add_end_function()
    # the same loop body on a cooperative thread group, a branching select of two variables, the single-timing copies
    g.gen_add_code_line("template <typename T>")
    g.gen_add_code_line("__global__ void user_kernel_tgrp(T *d_y, const T *d_x) {", True)
    g.gen_add_code_line("cgrps::thread_group tgrp = cgrps::this_thread_block();")
    g.gen_add_code_line("__shared__ T s_x[12]; __shared__ T s_y[12];")
    g.gen_kernel_load_inputs_single_timing("x", "12", use_thread_group=True)
    g.gen_add_parallel_loop("ind", "12", use_thread_group=True)
    g.gen_add_multi_threaded_select("ind", "<", ["4", "8", "12"], [("T", "a", ["T(1)", "T(2)", "T(3)"]), ("int", "src", ["0", "4", "8"])])
    g.gen_add_code_line(f"if {g.gen_var_in_list('ind', ['0', '11'])} {{ a = T(0); }}")
    g.gen_add_code_line("s_y[ind] = a * s_x[src];")
    g.gen_add_end_control_flow()
    g.gen_add_debug_print_code_line('if (false) printf("%f\\n", (double)s_y[0]);', use_thread_group=True)
    g.gen_kernel_save_result_single_timing("y", "12", use_thread_group=True)
    g.gen_add_end_function()
    # motion cross-product helpers through gen_mx_func_call_for_cpp: fixed S (joint 0 alone) and run-time S (all joints)
    g.gen_add_code_line("template <typename T>")
    g.gen_add_code_line("__global__ void user_kernel_mx(T *d_y, const T *d_x, const int S_ind) {", True)
    g.gen_add_code_line("__shared__ T s_src[6]; __shared__ T s_dst[24];")
    g.gen_add_parallel_loop("ind", "6")
    g.gen_add_code_line("s_src[ind] = d_x[ind];")
    g.gen_add_end_control_flow()
    g.gen_add_parallel_loop("ind", "24")
    g.gen_add_code_line("s_dst[ind] = static_cast<T>(1);")
    g.gen_add_end_control_flow()
    g.gen_add_sync()
    g.gen_add_serial_ops()
    g.gen_add_code_line("const T s_scale = static_cast<T>(0.5);")
    g.gen_mx_func_call_for_cpp(inds=[0])
    g.gen_mx_func_call_for_cpp(inds=[0], PEQ_FLAG=True, updated_var_names=dict(s_dst_name="&s_dst[6]"))
    g.gen_mx_func_call_for_cpp(SCALE_FLAG=True, updated_var_names=dict(s_dst_name="&s_dst[12]"))
    g.gen_mx_func_call_for_cpp(PEQ_FLAG=True, SCALE_FLAG=True, updated_var_names=dict(s_dst_name="&s_dst[18]"))
    g.gen_add_end_control_flow()
    g.gen_add_sync()
    g.gen_kernel_save_result_single_timing("y", "24", load_from_name="s_dst")
    g.gen_add_end_function()
    assert g.indent_level == 0
    return g.code_str


HARNESS = r'''
#include <vector>
int main() {
    const int N = 37;
    std::vector<float> x(N * 12), y(N * 12), z(N), y2(12);
    for (int i = 0; i < N * 12; i++) x[i] = 0.25f * (i % 17) - 1.0f;
    float *dx, *dy, *dz, *dy2;
    cudaMalloc(&dx, x.size() * 4); cudaMalloc(&dy, y.size() * 4); cudaMalloc(&dz, z.size() * 4); cudaMalloc(&dy2, 48);
    cudaMemcpy(dx, x.data(), x.size() * 4, cudaMemcpyHostToDevice);
    user_kernel<float><<<dim3(5, 2), dim3(8, 2)>>>(dy, dz, dx, N);
    user_kernel_tgrp<float><<<1, 32>>>(dy2, dx);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("launch failed\n"); return 2; }
    cudaMemcpy(y.data(), dy, y.size() * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(z.data(), dz, z.size() * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(y2.data(), dy2, 48, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int k = 0; k < N; k++) {
        float acc = 0;
        for (int i = 0; i < 12; i++) { acc += x[k * 12 + i]; bad += y[k * 12 + i] != (i < 6 ? 2.0f : 3.0f) * x[k * 12 + i]; }
        bad += z[k] != acc;
    }
    for (int i = 0; i < 12; i++) {
        const float a = (i == 0 || i == 11) ? 0.0f : (i < 4 ? 1.0f : (i < 8 ? 2.0f : 3.0f));
        bad += y2[i] != a * x[i < 4 ? 0 : (i < 8 ? 4 : 8)];
    }
    {   // user_kernel_mx: [mxS(x), 1 + mxS(x), 0.5 mxS2(x), 1 + 0.5 mxS2(x)] with S = branched6's joint-0 axis and S2 = 4
        std::vector<float> ym(24);
        float* dym; cudaMalloc(&dym, 96);
        user_kernel_mx<float><<<1, 32>>>(dym, dx, 4);
        cudaDeviceSynchronize();
        cudaMemcpy(ym.data(), dym, 96, cudaMemcpyDeviceToHost);
        auto crm = [&](int r, int c) {   // entry (r, c) of the motion cross-product matrix of x[0..5]
            const float* v = x.data();
            const float w[3] = {v[0], v[1], v[2]}, l[3] = {v[3], v[4], v[5]};
            auto skew = [](const float* a, int i, int j) { if (i == j) return 0.0f; const int k = 3 - i - j; const float s = ((j - i + 3) % 3 == 1) ? -1.0f : 1.0f; return s * a[k]; };
            if (r < 3 && c < 3) return skew(w, r, c);
            if (r >= 3 && c < 3) return skew(l, r - 3, c);
            if (r >= 3 && c >= 3) return skew(w, r - 3, c - 3);
            return 0.0f;
        };
        const int S0 = MX_FIXED_S;
        for (int i = 0; i < 6; i++) {
            bad += ym[i] != crm(i, S0);
            bad += ym[6 + i] != 1.0f + crm(i, S0);
            bad += ym[12 + i] != 0.5f * crm(i, 4);
            bad += ym[18 + i] != 1.0f + 0.5f * crm(i, 4);
        }
    }
    printf("bad %d\n", bad);
    return bad != 0;
}
'''


def test_helper_text_and_expression_builders():
    g = GRiDCodeGenerator(load_robot("branched6"))
    g.code_str, g.indent_level = "", 0
    g.gen_add_parallel_loop("i", "n")
    g.gen_add_serial_ops()
    g.gen_add_end_control_flow()
    g.gen_add_end_control_flow()
    assert g.indent_level == 0
    lines = g.code_str.splitlines()
    assert lines[0] == "for(int i = threadIdx.x + threadIdx.y*blockDim.x; i < n; i += blockDim.x*blockDim.y){"
    assert lines[1].strip() == "if(threadIdx.x == 0 && threadIdx.y == 0){" and lines[1].startswith("    ")
    assert g.gen_var_in_list("j", ["1"]) == "(j == 1)" and g.gen_var_in_list("j", ["1", "2"]) == "((j == 1) || (j == 2))"
    assert g.gen_var_not_in_list("j", ["1", "2"]) == "((j != 1) && (j != 2))"
    assert g.gen_static_array_ind_2d(2, 3) == 15 and g.gen_static_array_ind_3d(1, 2, 3) == 51
    g.code_str = ""
    g.gen_add_multi_threaded_select("ind", "<", ["6", "12"], [("int", "jid", ["0", "1"])])
    assert g.code_str.splitlines()[1] == "int jid = (ind < 6) * 0 + (ind >= 6) * 1;"
    g.code_str = ""
    g.gen_add_multi_threaded_select("ind", "==", ["0", "1", "2"], [(None, "p", ["a", "b", "c"])], USE_NON_BRANCH_ALWAYS=True)
    assert g.code_str.splitlines()[1] == "p = (ind == 0) * a + (ind == 1) * b + (ind == 2) * c;"
    with pytest.raises(NotImplementedError):
        g.gen_add_parallel_loop("i", "n", use_thread_group=True, block_level=True)


@pytest.mark.parametrize("name", ["iiwa14", "hyq", "branched6"])
def test_topology_sparsity_helpers(name):
    """Against the definition (reference helpers/_topology_helpers.py:515-542) and, in the build container, against the
    reference's own function on the same robot object."""
    r = load_robot(name)
    g = GRiDCodeGenerator(r)
    n = r.get_num_joints()
    dva_pp, dva_pj, run_dva, df_pp, df_pj, run_df, df_self = g.gen_topology_sparsity_helpers_python()
    assert dva_pp == sum(len(r.get_ancestors_by_id(j)) + 1 for j in range(n)) == sum(dva_pj)
    assert df_pp == sum(len(r.get_ancestors_by_id(j)) + len(r.get_subtree_by_id(j)) for j in range(n)) == sum(df_pj)
    assert run_dva[0] == 0 and run_dva[n] == dva_pp and all(run_dva[j + 1] - run_dva[j] == dva_pj[j] for j in range(n))
    assert run_df[0] == 0 and all(run_df[j + 1] - run_df[j] == df_pj[j] for j in range(n - 1))
    assert df_self == [len(r.get_ancestors_by_id(j)) for j in range(n)]
    init = g.gen_topology_sparsity_helpers_python(INIT_MODE=True)
    assert init[0] == [str(v) for v in df_self] and len(init[2]) == n + 1
    if os.path.isdir("/root/reference"):
        import sys
        sys.path.insert(0, "/root")
        from reference import GRiDCodeGenerator as RefGen
        ref = RefGen(r)
        assert tuple(ref.gen_topology_sparsity_helpers_python()) == (dva_pp, dva_pj, run_dva, df_pp, df_pj, run_df, df_self)
        assert tuple(ref.gen_topology_sparsity_helpers_python(INIT_MODE=True)) == tuple(init)


def _fixed_S():
    return int(list(load_robot("branched6").get_S_by_id(0)).index(1))


def test_mx_call_generator_text():
    r = load_robot("branched6")
    g = GRiDCodeGenerator(r)
    g.code_str, g.indent_level = "", 0
    g.gen_mx_func_call_for_cpp(inds=[0])
    S0 = _fixed_S()
    assert g.code_str.strip() == f"mx{S0}<T>(s_dst, s_src);"
    g.code_str = ""
    mixed = not r.are_Ss_identical(list(range(r.get_num_pos())))
    g.gen_mx_func_call_for_cpp(PEQ_FLAG=True, SCALE_FLAG=True, updated_var_names=dict(s_scale_name="qd_j"))
    want = "mxX_peq_scaled<T>(s_dst, s_src, qd_j, S_ind);" if mixed else f"mx{S0}_peq_scaled<T>(s_dst, s_src, qd_j);"
    assert g.code_str.strip() == want


def _nvcc():
    from gridcodegenerator_b200.runtime import _nvcc
    return _nvcc()


def test_user_kernel_compiles():
    src = _user_kernel_source()
    with tempfile.TemporaryDirectory() as d:
        cu = os.path.join(d, "user.cu")
        with open(cu, "w") as f:
            f.write(src + HARNESS)
        res = subprocess.run([_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", f"-DMX_FIXED_S={_fixed_S()}", "-c", cu, "-o", os.path.join(d, "user.o")],
                             stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        assert res.returncode == 0, res.stdout[-3000:]


@pytest.mark.gpu
def test_user_kernel_runs():
    src = _user_kernel_source()
    with tempfile.TemporaryDirectory() as d:
        cu, exe = os.path.join(d, "user.cu"), os.path.join(d, "user")
        with open(cu, "w") as f:
            f.write(src + HARNESS)
        subprocess.check_call([_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", f"-DMX_FIXED_S={_fixed_S()}", cu, "-o", exe])
        res = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
        assert res.returncode == 0 and "bad 0" in res.stdout, res.stdout
