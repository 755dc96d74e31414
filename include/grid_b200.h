/*
 * grid_b200.h -- C ABI of one robot-specialised GRiD library (libgrid_<robot>.so).
 *
 * One shared library per robot, like the reference's one grid.cuh per robot
 * (reference README.md:10-15, GRiDCodeGenerator.py:434-437).  Each entry point replaces the host tier
 * of one algorithm of the reference's emitted header; the reference function it stands in for is cited
 * per declaration (files relative to /root/reference).
 *
 * Conventions
 *   - all matrices column-major, batches knot-major (reference helpers/_code_generation_helpers.py:57-61,
 *     :147-164): knot k's inputs start at in + k*stride.
 *   - `d_*` pointers are device memory owned by the caller; calls are asynchronous on `stream`
 *     (a cudaStream_t passed as void*; NULL = default stream) and never allocate after the first call
 *     for a robot/dtype on a device.  `h_*` entry points take host buffers (pinned or pageable), run
 *     H2D -> kernel -> D2H internally and return when the result is in h_out.
 *   - every function returns 0 on success or a cudaError_t value; grid_last_error() describes it.
 *     Nothing calls exit() (the reference's gpuErrchk does: GRiDCodeGenerator.py:279-286).
 *   - the current CUDA device of the calling thread is used (there is no `device` argument: like the CUDA runtime
 *     itself the library is device-implicit, and its per-device state -- robot-model handles per (device, stream), host
 *     workspace per device -- is keyed on cudaGetDevice(); one process may drive several GPUs from several threads).
 *   - gravity is the positive magnitude (9.81), as in the reference kernels
 *     (algorithms/_inverse_dynamics.py:123) == GRAVITY = -9.81 of _test.py:13-14.
 */
#ifndef GRID_B200_H
#define GRID_B200_H

#ifdef __cplusplus
extern "C" {
#endif

/* ---- introspection (reference constants NUM_JOINTS etc., GRiDCodeGenerator.py:86-111) ---- */
int grid_num_joints(void);
int grid_num_ees(void);           /* number of leaf links = end effectors (reference NUM_EES, GRiDCodeGenerator.py:96) */
const char *grid_robot_name(void);
const char *grid_last_error(void);
int grid_abi_version(void);
/* shared-memory bytes per CTA, threads per CTA, per-thread state elements of an algorithm's kernel
 * algo: 0 ID, 1 ID(qdd), 2 MINV, 3 FD, 4 ID_DU, 5 ID_DU(qdd), 6 FD_DU, 7 FD_DU(qdd,Minv), 8 ABA, 9 CRBA,
 *       10 IDSVA_SO, 11 FDSVA_SO (-1 when the library was built without second-order kernels),
 *       12 EE pose, 13 EE pose gradient, 14 EE pose Hessian;
 * is_f64: 0/1
 * what: 0 smem bytes, 1 threads, 2 state slots, 3 state in shared memory (0/1), 4 TMA-staged I/O (0/1),
 *       5 persistent grid size for num_knots = 1<<20 on the current device */
long grid_kernel_info(int algo, int is_f64, int what);

/* ---- inverse dynamics c = RNEA(q, qd[, qdd])
 * replaces inverse_dynamics<T,USE_QDD_FLAG,USE_COMPRESSED_MEM> (algorithms/_inverse_dynamics.py:440-512)
 * d_q_qd: [q|qd] per knot, `stride` elements apart (2n compressed, 3n for a q_qd_u buffer);
 * d_qdd: n per knot, dense, or NULL for qdd = 0; d_c: n per knot. */
int grid_inverse_dynamics_f32(const float *d_q_qd, int stride, const float *d_qdd, float *d_c, int num_knots,
                              float gravity, void *stream);
int grid_inverse_dynamics_f64(const double *d_q_qd, int stride, const double *d_qdd, double *d_c, int num_knots,
                              double gravity, void *stream);

/* ---- RNEA's per-joint spatial velocity, acceleration and (accumulated) force
 * replaces inverse_dynamics_inner_vaf / inverse_dynamics_vaf_device (algorithms/_inverse_dynamics.py:33-321; banner
 * GRiDCodeGenerator.py:326-329): the reference's s_vaf array per knot, [v (6n) | a (6n) | f (6n)], the 6-vector
 * ([angular; linear]) of joint j at offset 6 j of each block.  Inputs as grid_inverse_dynamics; d_vaf: 18n per knot. */
int grid_inverse_dynamics_vaf_f32(const float *d_q_qd, int stride, const float *d_qdd, float *d_vaf, int num_knots,
                                  float gravity, void *stream);
int grid_inverse_dynamics_vaf_f64(const double *d_q_qd, int stride, const double *d_qdd, double *d_vaf, int num_knots,
                                  double gravity, void *stream);

/* ---- direct inverse mass matrix, dense symmetric n x n per knot
 * replaces direct_minv<T,USE_COMPRESSED_MEM> (algorithms/_direct_minv.py:527-588) */
int grid_direct_minv_f32(const float *d_q, int stride, float *d_Minv, int num_knots, void *stream);
int grid_direct_minv_f64(const double *d_q, int stride, double *d_Minv, int num_knots, void *stream);

/* ---- forward dynamics qdd = Minv (u - c)
 * replaces forward_dynamics<T> (algorithms/_forward_dynamics.py:199-260); d_q_qd_u: [q|qd|u] */
int grid_forward_dynamics_f32(const float *d_q_qd_u, int stride, float *d_qdd, int num_knots, float gravity,
                              void *stream);
int grid_forward_dynamics_f64(const double *d_q_qd_u, int stride, double *d_qdd, int num_knots, double gravity,
                              void *stream);

/* ---- gradient of inverse dynamics [dc/dq | dc/dqd], n x 2n per knot
 * replaces inverse_dynamics_gradient<T,USE_QDD_FLAG,USE_COMPRESSED_MEM>
 * (algorithms/_inverse_dynamics_gradient.py:890-960) */
int grid_inverse_dynamics_gradient_f32(const float *d_q_qd, int stride, const float *d_qdd, float *d_dc_du,
                                       int num_knots, float gravity, void *stream);
int grid_inverse_dynamics_gradient_f64(const double *d_q_qd, int stride, const double *d_qdd, double *d_dc_du,
                                       int num_knots, double gravity, void *stream);

/* ---- gradient of forward dynamics [dqdd/dq | dqdd/dqd] = -Minv [dc/dq | dc/dqd], n x 2n per knot
 * replaces forward_dynamics_gradient<T,false> (algorithms/_forward_dynamics_gradient.py:186-249) */
int grid_forward_dynamics_gradient_f32(const float *d_q_qd_u, int stride, float *d_df_du, int num_knots,
                                       float gravity, void *stream);
int grid_forward_dynamics_gradient_f64(const double *d_q_qd_u, int stride, double *d_df_du, int num_knots,
                                       double gravity, void *stream);
/* replaces forward_dynamics_gradient<T,true> (qdd and Minv supplied; same file :64-105, :223-228) */
int grid_forward_dynamics_gradient_qdd_minv_f32(const float *d_q_qd, int stride, const float *d_qdd,
                                                const float *d_Minv, float *d_df_du, int num_knots,
                                                float gravity, void *stream);
int grid_forward_dynamics_gradient_qdd_minv_f64(const double *d_q_qd, int stride, const double *d_qdd,
                                                const double *d_Minv, double *d_df_du, int num_knots,
                                                double gravity, void *stream);

/* ---- grad FD together with its by-products, one launch: per knot 3 n^2 + n elements
 *   d_fd_lin = [dqdd/dq (n x n) | dqdd/dqd (n x n) | Minv = dqdd/du (n x n, dense symmetric) | qdd (n)], column-major
 * i.e. the (A, B, f) of the linearised dynamics an MPC / DDP caller needs.  The reference attempts the Minv side
 * output inside its grad FD body without a barrier (algorithms/_forward_dynamics_gradient.py:56-59, racy) and
 * otherwise needs a second direct_minv / forward_dynamics launch (host tiers _direct_minv.py:455-560,
 * _forward_dynamics.py:117-230); this entry point replaces that launch sequence. */
int grid_forward_dynamics_linearization_f32(const float *d_q_qd_u, int stride, float *d_fd_lin, int num_knots,
                                            float gravity, void *stream);
int grid_forward_dynamics_linearization_f64(const double *d_q_qd_u, int stride, double *d_fd_lin, int num_knots,
                                            double gravity, void *stream);

/* ---- forward dynamics by the articulated-body algorithm, qdd n per knot; d_q_qd_tau: [q|qd|tau]
 * replaces aba<T> (algorithms/_aba.py:539-596; kernel :482-537) */
int grid_aba_f32(const float *d_q_qd_tau, int stride, float *d_qdd, int num_knots, float gravity, void *stream);
int grid_aba_f64(const double *d_q_qd_tau, int stride, double *d_qdd, int num_knots, double gravity, void *stream);

/* ---- joint-space inertia matrix M(q) by the composite-rigid-body algorithm, dense symmetric n x n per knot
 * replaces crba<T,USE_COMPRESSED_MEM> (algorithms/_crba.py:310-375; kernel :254-308); only q (the first n
 * elements of each knot's row) is read */
int grid_crba_f32(const float *d_q, int stride, float *d_M, int num_knots, void *stream);
int grid_crba_f64(const double *d_q, int stride, double *d_M, int num_knots, void *stream);

/* ---- second-order tensors, 4 n^3 per knot, each n x n x n with index i*n*n + j*n + k
 * grid_has_second_order(): 1 if this robot's library contains them (robots up to 8 joints by default), else the
 * two entry points return cudaErrorNotSupported.
 * idsva_so replaces idsva_so_host<T> (algorithms/_idsva_so.py:1030-1093; kernel :958-1028); d_q_qd_qdd: [q|qd|qdd]
 *   [0] d2tau_i/dq_j dq_k   [1] d2tau_i/dqd_j dqd_k   [2] d2tau_i/dqd_j dq_k   [3] dM_ij/dq_k
 * fdsva_so replaces fdsva_so<T> (algorithms/_fdsva_so.py:246-307; kernel :161-244); d_q_qd_u: [q|qd|tau]
 *   [0] d2qdd_i/dq_j dq_k   [1] d2qdd_i/dqd_j dqd_k   [2] d2qdd_i/dqd_j dq_k   [3] d2qdd_i/dtau_j dq_k */
int grid_has_second_order(void);
int grid_idsva_so_f32(const float *d_q_qd_qdd, int stride, float *d_idsva_so, int num_knots, float gravity, void *stream);
int grid_idsva_so_f64(const double *d_q_qd_qdd, int stride, double *d_idsva_so, int num_knots, double gravity, void *stream);
int grid_fdsva_so_f32(const float *d_q_qd_u, int stride, float *d_df2, int num_knots, float gravity, void *stream);
int grid_fdsva_so_f64(const double *d_q_qd_u, int stride, double *d_df2, int num_knots, double gravity, void *stream);

/* ---- end-effector (leaf link frame) poses [x y z roll pitch yaw] in the base frame, their gradient and Hessian
 * with respect to q; only q (first n elements of each knot's row) is read.
 * replace end_effector_pose<T,..>, end_effector_pose_gradient<T,..>, end_effector_pose_gradient_hessian<T,..>
 * (algorithms/_eepose_gradient_hessian.py:253-310, :632-690, :1157-1220; kernels :197-251, :576-630, :1097-1155)
 *   d_eePos  [6*ee + o]                      6 NUM_EES per knot
 *   d_deePos [(ee*n + j)*6 + o]              d pose_o / dq_j          (reference s_deePos[6*deeInd + outputInd])
 *   d_d2eePos[(ee*6 + o)*n*n + i*n + j]      d2 pose_o / dq_i dq_j    (reference s_d2eePos[outputInd*n*n + i*n + j])
 * gradient_hessian fills both outputs like the reference kernel (d_deePos may be NULL to skip the gradient). */
int grid_has_ee_hessian(void);   /* 0: robot above 16 joints, gradient_hessian returns cudaErrorNotSupported */
int grid_end_effector_pose_f32(const float *d_q, int stride, float *d_eePos, int num_knots, void *stream);
int grid_end_effector_pose_f64(const double *d_q, int stride, double *d_eePos, int num_knots, void *stream);
int grid_end_effector_pose_gradient_f32(const float *d_q, int stride, float *d_deePos, int num_knots, void *stream);
int grid_end_effector_pose_gradient_f64(const double *d_q, int stride, double *d_deePos, int num_knots, void *stream);
int grid_end_effector_pose_gradient_hessian_f32(const float *d_q, int stride, float *d_deePos, float *d_d2eePos,
                                                int num_knots, void *stream);
int grid_end_effector_pose_gradient_hessian_f64(const double *d_q, int stride, double *d_deePos, double *d_d2eePos,
                                                int num_knots, void *stream);

/* ---- host-buffer entry points (the reference's default host semantics: H2D, kernel, D2H, blocking;
 * algorithms/_forward_dynamics_gradient.py:217-245).  Dense strides (3n in, 2n^2 / 3n^2+n / n out).
 * Pinned, device-mapped caller buffers (cudaHostAlloc / cudaHostRegister) can take a zero-copy transport with
 * GRID_HOST_ZEROCOPY=1 in the environment: one kernel reads and writes the host buffers directly over PCIe;
 * GRID_HOST_CHUNK sets the pipeline's chunk size in knot points (default 8192). */
int grid_forward_dynamics_gradient_host_f32(const float *h_q_qd_u, float *h_df_du, int num_knots, float gravity);
int grid_forward_dynamics_gradient_host_f64(const double *h_q_qd_u, double *h_df_du, int num_knots, double gravity);
int grid_forward_dynamics_linearization_host_f32(const float *h_q_qd_u, float *h_fd_lin, int num_knots, float gravity);
int grid_forward_dynamics_linearization_host_f64(const double *h_q_qd_u, double *h_fd_lin, int num_knots, double gravity);
int grid_forward_dynamics_host_f32(const float *h_q_qd_u, float *h_qdd, int num_knots, float gravity);
int grid_inverse_dynamics_host_f32(const float *h_q_qd_u, float *h_c, int num_knots, float gravity);
/* release the cached device workspace of the host entry points and the robot model handles */
int grid_release(void);

/* ---- host-buffer variants of the remaining algorithms (the reference host functions' default mode: H2D -> kernel ->
 * D2H, blocking; same chunked three-stream pipeline as grid_forward_dynamics_gradient_host_*).  Dense rows:
 * h_q n, h_q_qd 2n, h_q_qd_tau / h_q_qd_qdd / h_q_qd_u 3n per knot; h_qdd (n per knot) may be NULL for qdd = 0. */
int grid_direct_minv_host_f32(const float *h_q, float *h_Minv, int num_knots);
int grid_direct_minv_host_f64(const double *h_q, double *h_Minv, int num_knots);
int grid_inverse_dynamics_gradient_host_f32(const float *h_q_qd, const float *h_qdd, float *h_dc_du, int num_knots, float gravity);
int grid_inverse_dynamics_gradient_host_f64(const double *h_q_qd, const double *h_qdd, double *h_dc_du, int num_knots, double gravity);
int grid_aba_host_f32(const float *h_q_qd_tau, float *h_qdd, int num_knots, float gravity);
int grid_aba_host_f64(const double *h_q_qd_tau, double *h_qdd, int num_knots, double gravity);
int grid_crba_host_f32(const float *h_q, float *h_M, int num_knots);
int grid_crba_host_f64(const double *h_q, double *h_M, int num_knots);
int grid_idsva_so_host_f32(const float *h_q_qd_qdd, float *h_out, int num_knots, float gravity);
int grid_idsva_so_host_f64(const double *h_q_qd_qdd, double *h_out, int num_knots, double gravity);
int grid_fdsva_so_host_f32(const float *h_q_qd_u, float *h_out, int num_knots, float gravity);
int grid_fdsva_so_host_f64(const double *h_q_qd_u, double *h_out, int num_knots, double gravity);

/* ---- cumulative joint transforms: out[16*j + r + 4*c] = (base <- frame of joint j)[r][c], 4 x 4 column-major per joint
 * replaces what X_single_thread / X_warp leave in s_jointXforms (algorithms/_eepose_gradient_hessian.py:1226-1398; the
 * emitted header also provides those two device functions with the reference's signatures). d_q: rows starting with q. */
int grid_joint_transforms_f32(const float *d_q, int stride, float *d_jointXforms, int num_knots, void *stream);
int grid_joint_transforms_f64(const double *d_q, int stride, double *d_jointXforms, int num_knots, void *stream);

/* ---- end-effector functions of ONE fixed kinematic target (a fixed joint chosen when the library is generated)
 * replaces end_effector_pose_<target> / _gradient_<target> / _gradient_hessian_<target> that the reference emits for
 * gen_all_code(fixed_target_name=...) (GRiDCodeGenerator.py:309-310, :417; algorithms/_eepose_gradient_hessian.py:31-38).
 * grid_fixed_target_name() is "" when the library has none (the calls then return cudaErrorNotSupported).
 * pose: 6 per knot; gradient: out[j*6 + o]; Hessian: out[o*n*n + i*n + j]; d_grad may be NULL. */
const char *grid_fixed_target_name(void);
int grid_fixed_target_pose_f32(const float *d_q, int stride, float *d_pose, int num_knots, void *stream);
int grid_fixed_target_pose_f64(const double *d_q, int stride, double *d_pose, int num_knots, void *stream);
int grid_fixed_target_pose_gradient_f32(const float *d_q, int stride, float *d_grad, int num_knots, void *stream);
int grid_fixed_target_pose_gradient_f64(const double *d_q, int stride, double *d_grad, int num_knots, void *stream);
int grid_fixed_target_pose_gradient_hessian_f32(const float *d_q, int stride, float *d_grad, float *d_hess, int num_knots,
                                                void *stream);
int grid_fixed_target_pose_gradient_hessian_f64(const double *d_q, int stride, double *d_grad, double *d_hess, int num_knots,
                                                void *stream);

/* ---- grad FD without any allocation
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same inputs / outputs / arithmetic as grid_forward_dynamics_gradient_f32 (one-lane-per-knot kernel form), but the
 * kernel's global scratch is a file-scope static __device__ array of the library (csrc/grid_noalloc.cu) instead of a
 * cudaMalloc'ed robot-model handle: no cudaMalloc / cudaFree / allocating library call at set-up or per call.
 *   - limit: num_knots <= GRID_NOALLOC_MAX_KNOTS; larger batches return cudaErrorInvalidValue (see grid_last_error()).
 *   - cost: every GPU that loads the library holds grid_noalloc_scratch_bytes() of device memory until exit
 *     (HyQ 7 MB, Atlas 17 MB, robots whose grad FD needs no scratch 256 B); at most 32 CTAs run.
 *   - every call on a device shares that one array (it keeps the previous call's contents; nothing in it has to start
 *     at zero): calls that OVERLAP on the GPU -- two streams, two host threads on one device -- race; serialise them. */
#define GRID_NOALLOC_MAX_KNOTS 8192
int grid_forward_dynamics_gradient_noalloc_f32(const float *d_q_qd_u, int stride, float *d_df_du, int num_knots,
                                               float gravity, void *stream);
int grid_noalloc_max_knots(void);
long grid_noalloc_scratch_bytes(void);

#ifdef __cplusplus
}
#endif
#endif /* GRID_B200_H */
