/*
 * mpcgpu_b200.h -- C ABI of the B200-native (sm_100a) linear-system hot path of MPCGPU:
 *     form_schur_system  (Schur complement S, symmetric-stair preconditioner Pinv, gamma, G^-1)
 *     pcg                (block-tridiagonal preconditioned conjugate gradient)
 * plus the fused KKT solve (S / Pinv never materialised in HBM).
 *
 * Drop-in boundary: every entry point cites the reference interface it replaces (paths
 * relative to the MPCGPU checkout).  Plain pointers and sizes only; no CUDA or torch types
 * in the signatures (a stream is passed as `void*` holding a cudaStream_t, NULL = legacy
 * default stream, which is what the reference launches on: include/pcg/sqp.cuh:275,
 * include/pcg/linsys_setup.cuh:656).
 *
 * Conventions inherited from the reference:
 *   - caller owns every device buffer; nothing is allocated per call;
 *   - all blocks are COLUMN-MAJOR; buffer sizes as include/pcg/sqp.cuh:54-58,111,130:
 *       G     (s*s+c*c)*N - c*c      [Q_k | R_k] per knot (last knot Q only)   IN/OUT (G -> G^-1)
 *       C     (s*s+s*c)*(N-1)        [Abar_k | Bbar_k], Abar = -A, Bbar = -B
 *       g     (s+c)*N - c            [q_k | r_k]
 *       c     s*N                    constraint residuals (c[0:s] never read)
 *       S,Pinv 3*s*s*N               [N][3][s*s] block rows (left, diag, right)
 *       gamma, lambda  s*N
 *   - batched calls: instance b's buffer starts at base + b * <size above> (tight packing);
 *     batch == 1 is exactly the reference's single-instance layout;
 *   - d_lambda is IN (warm start, include/mpcsim.cuh:227,383) and OUT;
 *   - PCG non-convergence is data (max_iter_exit = 1), not an error
 *     (include/pcg/sqp.cuh:277,289);
 *   - return value: 0 on success, otherwise a cudaError_t value
 *     (cudaErrorInvalidValue for unsupported sizes); mpcgpu_b200_last_error() has the text.
 *   - supported sizes in this build: state_size 14, control_size 7 (Kuka IIWA,
 *     examples/track_iiwa_pcg.cu:16-17); fp32 (linsys_t, include/common/settings.cuh:49-57)
 *     with 1 <= knot_points <= 512, and fp64 (USE_DOUBLES 1) with 1 <= knot_points <= 256.
 * There is NO CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef MPCGPU_B200_H
#define MPCGPU_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MPCGPU_B200_ABI_VERSION 1

/* ABI version of the loaded library. */
int mpcgpu_b200_abi_version(void);

/* Text of the last error raised on the calling thread ("" if none). */
const char* mpcgpu_b200_last_error(void);

/* Bytes of device scratch `d_ws` the pcg / kkt_solve entry points want for these sizes.
 * Replaces the per-solve scratch the reference slot allocates (d_r, d_p, d_v_temp,
 * d_eta_new_temp: include/pcg/sqp.cuh:134-138).  May be 0 (cluster kernels keep all
 * iteration state on-chip); d_ws may then be NULL. */
size_t mpcgpu_b200_workspace_bytes(uint32_t state_size, uint32_t control_size,
                                   uint32_t knot_points, uint32_t batch);

/* form_schur_system<float>  (include/pcg/linsys_setup.cuh:620-657; kernel :565-612).
 * Writes S, Pinv (including zeros in the four corner blocks the reference leaves
 * unwritten), gamma, and overwrites G with G^-1 in place (linsys_setup.cuh:372-380). */
int mpcgpu_b200_schur_f32(uint32_t state_size, uint32_t control_size, uint32_t knot_points,
                          uint32_t batch, float* d_G, const float* d_C, const float* d_g,
                          const float* d_c, float rho, float* d_S, float* d_Pinv,
                          float* d_gamma, void* stream);

/* pcg<float, STATE_SIZE, KNOT_POINTS>  (GBD-PCG gpu_pcg.cuh; launch + argument list at
 * include/pcg/sqp.cuh:142-166,275; result read-back :276-277).
 * d_iters[b]: completed iterations; d_max_iter_exit[b]: 1 when max_iter was exhausted.
 * Exit test |r . Pinv r| < exit_tol (absolute). */
int mpcgpu_b200_pcg_f32(uint32_t state_size, uint32_t knot_points, uint32_t batch,
                        const float* d_S, const float* d_Pinv, const float* d_gamma,
                        float* d_lambda, uint32_t max_iter, float exit_tol, uint32_t* d_iters,
                        uint8_t* d_max_iter_exit, void* d_ws, void* stream);

/* Fused form_schur_system + pcg (include/pcg/sqp.cuh:243-277 as ONE launch): reads
 * G, C, g, c, lambda0; writes lambda, G^-1, iters, flags.  S and Pinv stay on-chip.
 * The iteration is the reference's two-reduction PCG.  Opt-in, read once per process: MPCGPU_B200_PCG_VARIANT=2 in the
 * environment makes the batched 64-knot tiling carry S p by a recurrence (4 % faster; same exit test and criteria, but
 * farther from the exact-arithmetic iterate at tight tolerances -- DESIGN.md 3).  No counterpart in the reference. */
int mpcgpu_b200_kkt_solve_f32(uint32_t state_size, uint32_t control_size, uint32_t knot_points,
                              uint32_t batch, float* d_G, const float* d_C, const float* d_g,
                              const float* d_c, float rho, float* d_lambda, uint32_t max_iter,
                              float exit_tol, uint32_t* d_iters, uint8_t* d_max_iter_exit,
                              void* d_ws, void* stream);

/* The same launch with the primal-step recovery as epilogue: compute_dz,
 * include/common/dz.cuh:124-133 (kernel :3-121), called at include/pcg/sqp.cuh:300-309.
 * d_dz: [(s+c)*N - c] per instance, the layout of d_xu.  Replaces one more launch and one more
 * pass over G^-1, C, g of the reference's SQP iteration. */
int mpcgpu_b200_kkt_solve_dz_f32(uint32_t state_size, uint32_t control_size, uint32_t knot_points,
                                 uint32_t batch, float* d_G, const float* d_C, const float* d_g,
                                 const float* d_c, float rho, float* d_lambda, float* d_dz,
                                 uint32_t max_iter, float exit_tol, uint32_t* d_iters,
                                 uint8_t* d_max_iter_exit, void* d_ws, void* stream);

/* The same with rho read from device memory (one float, shared by the batch) when the kernel
 * runs: every argument is then fixed, so the launch can be captured once into a CUDA graph and
 * replayed per SQP iteration while rho adapts (include/pcg/sqp.cuh:392-393,412-413). */
int mpcgpu_b200_kkt_solve_dz_dev_f32(uint32_t state_size, uint32_t control_size, uint32_t knot_points,
                                     uint32_t batch, float* d_G, const float* d_C, const float* d_g,
                                     const float* d_c, const float* d_rho, float* d_lambda,
                                     float* d_dz, uint32_t max_iter, float exit_tol,
                                     uint32_t* d_iters, uint8_t* d_max_iter_exit, void* d_ws,
                                     void* stream);

/* Same with one rho PER INSTANCE, d_rho[batch] in device memory: batched SQP iterations adapt every
 * instance's regularisation on their own (include/pcg/sqp.cuh:375-389), see shim/b200/sqp_batched.cuh. */
int mpcgpu_b200_kkt_solve_dz_rhovec_f32(uint32_t state_size, uint32_t control_size, uint32_t knot_points,
                                        uint32_t batch, float* d_G, const float* d_C, const float* d_g,
                                        const float* d_c, const float* d_rho, float* d_lambda,
                                        float* d_dz, uint32_t max_iter, float exit_tol,
                                        uint32_t* d_iters, uint8_t* d_max_iter_exit, void* d_ws,
                                        void* stream);

/* compute_dz alone (include/common/dz.cuh:124-133): d_Ginv = d_G after mpcgpu_b200_schur_f32 /
 * kkt_solve, d_C, d_g as given to them, d_lambda the multipliers.  Batched like the rest. */
int mpcgpu_b200_dz_f32(uint32_t state_size, uint32_t control_size, uint32_t knot_points,
                       uint32_t batch, const float* d_Ginv, const float* d_C, const float* d_g,
                       const float* d_lambda, float* d_dz, void* stream);

/* ---- fp64 twins (linsys_t = double when include/common/settings.cuh:49-57 is built with
 * USE_DOUBLES 1).  Same buffers, layouts, semantics and error behaviour as the _f32 entry points
 * above with every float replaced by double; the same cluster kernels instantiated for double
 * (16 knots per 128-thread CTA, S rows in registers, Pinv rows in shared memory, DFMA pairs).
 * Supported sizes: state_size 14, control_size 7, 1 <= knot_points <= 256 (16 CTAs per cluster). */
int mpcgpu_b200_schur_f64(uint32_t state_size, uint32_t control_size, uint32_t knot_points,
                          uint32_t batch, double* d_G, const double* d_C, const double* d_g,
                          const double* d_c, double rho, double* d_S, double* d_Pinv,
                          double* d_gamma, void* stream);
int mpcgpu_b200_pcg_f64(uint32_t state_size, uint32_t knot_points, uint32_t batch,
                        const double* d_S, const double* d_Pinv, const double* d_gamma,
                        double* d_lambda, uint32_t max_iter, double exit_tol, uint32_t* d_iters,
                        uint8_t* d_max_iter_exit, void* d_ws, void* stream);
int mpcgpu_b200_kkt_solve_f64(uint32_t state_size, uint32_t control_size, uint32_t knot_points,
                              uint32_t batch, double* d_G, const double* d_C, const double* d_g,
                              const double* d_c, double rho, double* d_lambda, uint32_t max_iter,
                              double exit_tol, uint32_t* d_iters, uint8_t* d_max_iter_exit,
                              void* d_ws, void* stream);
/* d_dz may be NULL (no compute_dz epilogue). */
int mpcgpu_b200_kkt_solve_dz_f64(uint32_t state_size, uint32_t control_size, uint32_t knot_points,
                                 uint32_t batch, double* d_G, const double* d_C, const double* d_g,
                                 const double* d_c, double rho, double* d_lambda, double* d_dz,
                                 uint32_t max_iter, double exit_tol, uint32_t* d_iters,
                                 uint8_t* d_max_iter_exit, void* d_ws, void* stream);
/* rho read from device memory (one double), cf. mpcgpu_b200_kkt_solve_dz_dev_f32 */
int mpcgpu_b200_kkt_solve_dz_dev_f64(uint32_t state_size, uint32_t control_size, uint32_t knot_points,
                                     uint32_t batch, double* d_G, const double* d_C, const double* d_g,
                                     const double* d_c, const double* d_rho, double* d_lambda,
                                     double* d_dz, uint32_t max_iter, double exit_tol, uint32_t* d_iters,
                                     uint8_t* d_max_iter_exit, void* d_ws, void* stream);
int mpcgpu_b200_dz_f64(uint32_t state_size, uint32_t control_size, uint32_t knot_points,
                       uint32_t batch, const double* d_Ginv, const double* d_C, const double* d_g,
                       const double* d_lambda, double* d_dz, void* stream);

/* Reference-STYLE PCG (one CTA per knot, 4 grid barriers per iteration, cooperative launch):
 * our restatement of the reference kernel's structure, kept as the in-repo
 * comparison kernel and as the kernel behind the header shim's pcg<T,S,N>.  batch == 1.
 * d_ws: at least (2*s*N + 2*N) floats (d_r, d_p, d_v_temp, d_eta_new_temp). */
int mpcgpu_b200_pcg_refstyle_f32(uint32_t state_size, uint32_t knot_points, const float* d_S,
                                 const float* d_Pinv, const float* d_gamma, float* d_lambda,
                                 uint32_t max_iter, float exit_tol, uint32_t* d_iters,
                                 uint8_t* d_max_iter_exit, void* d_ws, void* stream);

/* The same kernel for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The scratch (d_r, d_p, d_v_temp, d_eta_new_temp) lives in a file-scope device array of the library
 * (mpcgpu_b200/csrc/noalloc.cu) sized for state_size == 14 and knot_points <= mpcgpu_b200_noalloc_max_knots() == 512
 * (61 440 B, held by every GPU that loads the library until exit); other sizes are refused with
 * cudaErrorInvalidValue and a message.  The call itself allocates nothing (no cudaMalloc / cudaFree, no allocating
 * library routine).  All calls on a device share that array and it keeps the last call's contents (the kernel
 * writes every element before reading it, nothing needs clearing): calls that OVERLAP on the same GPU race --
 * keep them on one stream or serialise them. */
uint32_t mpcgpu_b200_noalloc_max_knots(void);
int mpcgpu_b200_pcg_refstyle_noalloc_f32(uint32_t state_size, uint32_t knot_points, const float* d_S,
                                         const float* d_Pinv, const float* d_gamma, float* d_lambda,
                                         uint32_t max_iter, float exit_tol, uint32_t* d_iters,
                                         uint8_t* d_max_iter_exit, void* stream);

/* ---- host-buffer solver object (the call a user of the reference's solver slot makes
 * when the KKT blocks live on the host, cf. the QDLDL slot's PCIe round trip
 * include/qdldl/sqp.cuh:268-273).  Owns persistent device buffers and streams; nothing is
 * allocated per solve. ---- */
typedef struct mpcgpu_b200_solver mpcgpu_b200_solver;

int mpcgpu_b200_solver_create(mpcgpu_b200_solver** out, int device, uint32_t state_size,
                              uint32_t control_size, uint32_t knot_points, uint32_t max_batch);
void mpcgpu_b200_solver_destroy(mpcgpu_b200_solver* solver);

/* Host buffers (pinned for full PCIe speed; pageable works).  h_lambda IN/OUT.
 * h_Ginv may be NULL (G^-1 not copied back).  Copies are pipelined with the kernels in
 * chunks; returns after all results are in the host buffers. */
int mpcgpu_b200_solver_solve_host_f32(mpcgpu_b200_solver* solver, uint32_t batch,
                                      const float* h_G, const float* h_C, const float* h_g,
                                      const float* h_c, float rho, float* h_lambda,
                                      uint32_t max_iter, float exit_tol, uint32_t* h_iters,
                                      uint8_t* h_max_iter_exit, float* h_Ginv);

/* Same with the warm start read from its own buffer: h_lambda0 (in) may differ from h_lambda (out);
 * h_lambda0 == NULL is the cold start lambda0 = 0 of the first SQP iteration
 * (include/mpcsim.cuh:201 zero-fills d_lambda): nothing is copied in for it. */
int mpcgpu_b200_solver_solve_host_io_f32(mpcgpu_b200_solver* solver, uint32_t batch,
                                         const float* h_G, const float* h_C, const float* h_g,
                                         const float* h_c, float rho, const float* h_lambda0,
                                         float* h_lambda, uint32_t max_iter, float exit_tol,
                                         uint32_t* h_iters, uint8_t* h_max_iter_exit, float* h_Ginv);

/* ---- long horizon split by knot range across the GPUs of one box (NEW: the reference has no
 * multi-GPU path; BASELINE.json configs[3]).  Rank r owns knots
 * [r*knots_local, (r+1)*knots_local) and holds only ITS slices of S / Pinv / gamma / lambda
 * (block rows of the reference layout, include/pcg/sqp.cuh:111,130).  Per PCG iteration the
 * boundary rows of S p and Pinv r and the two scalar sums travel as tagged 8-byte peer stores
 * over NVLink into a mailbox in each destination GPU; there is no NCCL call and no kernel
 * boundary inside the solve.  d_mailboxes: HOST array of `world` device pointers (this rank's
 * own mailbox and the peers' mailboxes mapped with mpcgpu_b200_ipc_open), each
 * mpcgpu_b200_knotsplit_mailbox_bytes() long and zeroed once.  `epoch` must be identical on
 * all ranks and grow from one solve to the next by an ODD amount larger than 2*max_iter+8
 * (mpcgpu_b200_knotsplit_epoch_stride(max_iter)): the tags of a solve are epoch + exchange counter, and
 * the parity of `epoch` selects which half of every mailbox the solve uses, so that a rank that has
 * already entered the next solve never overwrites a slot a slower rank has not read yet.
 * A poll that waits longer than MPCGPU_B200_KNOTSPLIT_TIMEOUT_MS (default 30000) for a peer records
 * {1, rank, expected tag, seen tag} for mpcgpu_b200_knotsplit_last_timeout() and traps (the launch
 * fails with a CUDA error on this rank instead of hanging the GPU).
 * ranks_here > 1 emulates several ranks in one cooperative launch on ONE GPU (tests). */
size_t mpcgpu_b200_knotsplit_mailbox_bytes(void);
uint32_t mpcgpu_b200_knotsplit_epoch_stride(uint32_t max_iter);
int mpcgpu_b200_knotsplit_last_timeout(uint32_t out[4]);   /* returns 1 if a poll of this process timed out */
/* Latency floor of one exchange: `reps` round trips of one tagged 8-byte peer store + local poll between two
 * ranks (me = 0 / 1) through the last slot of their mailboxes; *d_out_ns = device wall time of the reps. */
int mpcgpu_b200_knotsplit_pingpong(uint32_t me, void* d_my_mailbox, void* d_peer_mailbox, uint32_t reps,
                                   uint32_t tag0, unsigned long long* d_out_ns, void* stream);
/* a zeroed mailbox in its OWN cudaMalloc allocation (an IPC handle names a whole allocation) */
int mpcgpu_b200_knotsplit_mailbox_create(void** d_mailbox);
int mpcgpu_b200_knotsplit_mailbox_destroy(void* d_mailbox);
int mpcgpu_b200_knotsplit_pcg_f32(uint32_t state_size, uint32_t knots_local, uint32_t ranks_here,
                                  uint32_t rank_first, uint32_t world, uint32_t epoch,
                                  void* const* d_mailboxes, const float* d_S, const float* d_Pinv,
                                  const float* d_gamma, float* d_lambda, uint32_t max_iter,
                                  float exit_tol, uint32_t* d_iters, uint8_t* d_max_iter_exit,
                                  void* stream);
/* CUDA IPC plumbing for the mailboxes (one process per GPU): 64-byte handle out / in. */
int mpcgpu_b200_ipc_export(const void* d_ptr, void* handle64);
int mpcgpu_b200_ipc_open(const void* handle64, void** d_ptr);
int mpcgpu_b200_ipc_close(void* d_ptr);

/* Kernel launches issued by this library since load (all threads). */
uint64_t mpcgpu_b200_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* MPCGPU_B200_H */
