/* ofb200 - C ABI of the B200-native variational optical-flow hot path.
 *
 * The reference (kursawe/OpticalFlow) is pure Python and has no FFI of its own; the boundary it exposes for
 * this path is the Python function `variational_optical_flow` (source/optical_flow.py:715-1210).  Each entry
 * point below replaces one stage of that function and cites it.  The Python drop-in
 * (opticalflow_b200/optical_flow.py) binds these symbols with ctypes; INTEGRATION.md shows the stub a
 * maintainer of the reference would add.
 *
 * Conventions
 *   - return 0 on success, negative on error (text via ofb_last_error(), thread-local);
 *   - every data pointer is a DEVICE pointer owned by the caller (e.g. torch.Tensor.data_ptr());
 *     the library never allocates or frees device memory and keeps no global mutable state;
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream); calls are asynchronous on it,
 *     except ofb_solve, which synchronises the stream every `check_every` iterations to test convergence (the counter
 *     travels through mapped pinned memory, not the copy engine);
 *   - `dtype`: OFB_F32 or OFB_F64 - the arithmetic and storage type of all image/vector arguments;
 *   - B = number of frame pairs in the batch.  Images are [B][Nx][Ny] row-major; vectors are three SoA
 *     planes [B][3][Nx][Ny] in the order ux, uy, remodelling (the reference's interleaved numbering
 *     idx = 3*Ny*i + 3*j + q, :1291-1300, never exists on the device);
 *   - per-pair regularisation weights alpha (speed_alpha) and beta (remodelling_alpha) are device arrays
 *     of B doubles, so one batch can sweep regularisation weights (vary_regularisation, :1918-1998).
 */
#ifndef OFB200_H
#define OFB200_H
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OFB_F32 0
#define OFB_F64 1

#define OFB_PRECOND_BJACOBI 0   /* pointwise 3x3 block-Jacobi */
#define OFB_PRECOND_MG      1   /* geometric multigrid V-cycle, block-Jacobi smoother, dense coarsest solve */

#define OFB_FLAG_NO_STREAM 1    /* use the register-only sweep on every level (A/B measurements) */
#define OFB_FLAG_NO_FUSED_TAIL 2 /* run the small multigrid levels as separate launches (A/B measurements) */
#define OFB_FLAG_LEGACY_COARSE_BOUNDARY 4 /* node-mirror far boundary on every coarse level (A/B measurements) */

#define OFB_FLAG_TILE_FASTEST 16 /* streamed sweeps: neighbouring column tiles of a strip run concurrently (A/B measurements) */
#define OFB_FLAG_STRIP_ITEMS 64 /* streamed sweeps: fixed-size row strips dealt round-robin instead of the balanced walk (A/B) */
#define OFB_FLAG_NO_HALF_STEP_EXIT 128 /* BiCGStab: no convergence test after the first half step (A/B measurements) */
#define OFB_FLAG_PDL 32 /* programmatic dependent launch for the solver-loop kernels (measured neutral to -0.3%: off by default) */
#define OFB_FLAG_FUSED_PROLONG 512 /* float32 image level: prolongation inside the first post-smoothing sweep instead of its own pass (A/B) */
#define OFB_FLAG_SERIAL_SETUP 256 /* dense coarsest-level inverse on the solve's own stream instead of beside the first fine sweeps (A/B) */
#define OFB_FLAG_SLAB_NCCL_HALO 8 /* slab solves: halo rows by ncclSend/ncclRecv instead of peer stores (A/B measurements) */

#define OFB_FLAG_BF16_RESIDUAL 4096 /* float solves: the V-cycle's residual (read by the restriction only) as bfloat16 (measured slower: off by default) */
#define OFB_FLAG_FP32_COEF 2048 /* float solves: stream the coarse-level coefficient planes as float32 instead of their bfloat16 copies (A/B) */
#define OFB_FLAG_POLL_AHEAD 1024 /* ofb_solve: enqueue iteration k+1 before reading iteration k's convergence count (measured slower: off by default) */

#define OFB_ABI_VERSION 6

typedef struct ofb_solve_options {
    int    precond;        /* OFB_PRECOND_*            default OFB_PRECOND_MG */
    int    nu_pre;         /* pre-smoothing sweeps      default 1 */
    int    nu_post;        /* post-smoothing sweeps     default 2 */
    double omega;          /* Jacobi damping where no per-sweep value is given; default 0.8 */
    int    check_every;    /* convergence poll period in iterations (host sync); default 1 */
    int    rows_per_strip; /* 0 = choose from the batch size */
    int    coarse_sweeps;  /* Jacobi sweeps on the coarsest level when it is too large for the dense solve; default 30 */
    int    flags;          /* OFB_FLAG_* bits; default 0 */
    double omega_pre[4];   /* damping of pre-smoothing sweep s (0 = use omega); sweep 0 is the zero-guess sweep; default {1.0} */
    double omega_post[4];  /* damping of post-smoothing sweep s (0 = use omega); default {1.3, 0.5} */
    int    max_levels;     /* cap on the number of multigrid levels (0 = coarsen as far as possible) */
    int    reserved;
} ofb_solve_options;

int         ofb_version(void);
const char* ofb_last_error(void);
void        ofb_default_options(ofb_solve_options* opt);
/* number of CUDA kernels this library has launched in the calling process (diagnostics, bench.py) */
long long   ofb_launch_count(void);

/* number of pairs the calling thread's last ofb_solve / ofb_solve_slab left unconverged (-1: unknown); lets a caller
 * skip its fallback logic without reading converged_out back from the device */
int         ofb_last_unconverged(void);
/* number of pairs the calling thread's last ofb_solve / ofb_solve_slab restarted from their true residual (ofb_solve) */
int         ofb_last_restarts(void);

/* Bytes of device workspace ofb_solve needs for this shape (caller allocates, contents need not be zeroed). */
size_t ofb_workspace_bytes(int B, int Nx, int Ny, int dtype, int precond);

/* K1  Gaussian blur of F frames: replaces blur_movie (optical_flow.py:282-306 -> skimage.filters.gaussian,
 *     mode 'nearest', truncate 4).  `weights` = device array of 2*radius+1 doubles (normalised kernel),
 *     `tmp` = scratch of the same size as `in`.  in/out may not alias. */
int ofb_blur(const void* in, void* out, void* tmp, const double* weights, int radius,
             int F, int Nx, int Ny, int dtype, void* stream);

/* K2  right-hand side of the linear system: replaces the derivative block and the RHS stores
 *     (optical_flow.py:812-823, :889, :938, :962).  P = previous frames, C = current frames (both [B][Nx][Ny];
 *     for a movie they are two overlapping views of the same frame stack), b3 out. */
int ofb_build_rhs(const void* P, const void* C, void* b3, int B, int Nx, int Ny, int dtype, void* stream);

/* K3  y = A x with the reference's assembled matrix (optical_flow.py:836-1072), boundary rows included;
 *     matrix-free, nothing is assembled. */
int ofb_apply(const void* P, const void* x3, void* y3, const double* alpha, const double* beta,
              int B, int Nx, int Ny, int dtype, void* stream);

/* K3e y = S x, the same operator with the ghost ring eliminated (what the Krylov solver iterates on):
 *     ghost values of x are ignored (taken from the boundary rows), ghost rows of y are 0. */
int ofb_apply_eliminated(const void* P, const void* x3, void* y3, const double* alpha, const double* beta,
                         int B, int Nx, int Ny, int dtype, void* stream);

/* K5s one damped block-Jacobi sweep x_out = x_in + omega D^-1 (b - S x_in): the multigrid smoother and the
 *     dominant kernel of the solve (10 words/pixel); exposed for measurement and tests.  omega, rows_per_strip
 *     and flags are taken from `opt` (NULL = defaults). */
int ofb_smooth(const void* P, const void* b3, const void* x_in, void* x_out, const double* alpha, const double* beta,
               const ofb_solve_options* opt, int B, int Nx, int Ny, int dtype, void* stream);

/* K4/K5  solve A x = b: replaces the PETSc KSP solve (optical_flow.py:1081-1132; -ksp_type bcgs, rtol on the
 *     unpreconditioned residual, nonzero initial guess).  x3: in = initial guess, out = solution (ghost ring
 *     completed from the boundary rows).  Ghost-ring entries of b3 must be 0 (the reference never sets them,
 *     :837).  A pair whose b is all zero gets x = 0.
 *     Diagnostics (optical_flow.py:1135-1157), device arrays of length B, each may be NULL:
 *       relres_out      |b - A x| / |b| RECOMPUTED from the returned x in fp64 arithmetic - the independent residual the
 *                       reference prints at :1150-1151;
 *       recurrence_out  the BiCGStab recurrence residual the iteration stopped on (the KSP's own norm, :1140-1142);
 *       converged_out   1 when the true residual passes  |b - A x| <= rtol |b| + 8 u |D x|  (u = unit round-off of
 *                       `dtype`, D = diagonal of A; a row of A carries about four times its diagonal in absolute
 *                       off-diagonals, so 8 u |D x| ~ 2 u |A| |x|: the residual that storing x and the iterates in
 *                       `dtype` leaves behind whatever the solver does - negligible in float64, a few 1e-5 |b| in
 *                       float32 at alpha = 1e5, where the KSP's own test (the recurrence residual, which is what the
 *                       reference's `converged` reports, :1135-1138) still reaches rtol);
 *       iters_out       BiCGStab iterations.
 *     A pair whose recurrence converged (or broke down) while its true residual does not pass is restarted from
 *     r = b - A x, at most twice. */
int ofb_solve(const void* P, const void* b3, void* x3, const double* alpha, const double* beta,
              double rtol, int max_it, const ofb_solve_options* opt, void* workspace, size_t workspace_bytes,
              int* iters_out, double* relres_out, int* converged_out, double* recurrence_out,
              int B, int Nx, int Ny, int dtype, void* stream);

/* One application of the preconditioner z = M^-1 r (testing / diagnostics). */
int ofb_precondition(const void* P, const void* r3, void* z3, const double* alpha, const double* beta,
                     const ofb_solve_options* opt, void* workspace, size_t workspace_bytes,
                     int B, int Nx, int Ny, int dtype, void* stream);

/* K6  epilogue: replaces unpack + apply_constant_boundary_condition + functionals + scaling
 *     (optical_flow.py:1159-1191).  functionals: device array [B][3] doubles = L1, speed, remodelling
 *     functional of each pair (overwritten).  moments (may be NULL): device array [B][4] doubles = sum speed,
 *     sum speed^2, sum remodelling, sum remodelling^2 over all Nx*Ny pixels - the summary statistics
 *     vary_regularisation takes with np.mean/np.var (:1978-1981).  scale = delta_x/delta_t. */
int ofb_epilogue(const void* P, const void* C, const void* x3, void* vx, void* vy, void* speed, void* gamma,
                 double* functionals, double* moments, const double* alpha, const double* beta, double scale,
                 int B, int Nx, int Ny, int dtype, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Slab decomposition of ONE very large frame pair over the GPUs of a node.
 * The reference has no counterpart: its only scaling knob is host RAM (optical_flow.py:761).  One process per
 * GPU; each rank holds a slab of consecutive image rows and the ranks exchange one boundary row per field with
 * their neighbours before every stencil sweep (NCCL send/recv over NVLink), sum their partial dot products
 * (all-reduce of 4 doubles per pair) and all-gather the multigrid level `gather_level`, below which every rank
 * solves the whole coarse grid redundantly.
 *
 * Layout on rank r: local planes have owned+2 rows.  Row 0 is the last owned row of rank r-1 (the mirror ghost
 * row of the frame on rank 0), row owned+1 the first owned row of rank r+1 (ghost row on the last rank).
 * P (and C for the epilogue) must additionally be readable one row BEFORE row 0 and one row AFTER row owned+1
 * wherever that row is a neighbour's row: the level-1 coefficients of the first owned coarse row need the image
 * derivative on the halo row.  b3 must be 0 on rows 0 and owned+1 (ofb_build_rhs on the local slab does that).
 * ------------------------------------------------------------------------------------------------------- */
typedef struct ofb_slab {
    int rank, world;        /* this process, number of slabs */
    int Nx_global;          /* rows of the whole frame, ghost rows included */
    int rows_per_rank;      /* interior rows owned by every rank but the last  (ofb_slab_partition) */
    int gather_level;       /* first multigrid level replicated on every rank   (ofb_slab_partition) */
    int reserved;
    void* comm;             /* ofb_comm_create */
} ofb_slab;

/* Rows per rank (a multiple of 2^gather_level) and gather level for a frame; hint <= 0 = automatic. */
int ofb_slab_partition(int Nx_global, int Ny, int world, int gather_level_hint, int* rows_per_rank, int* gather_level);
/* First owned global row and number of owned rows of slab->rank (local Nx = owned + 2). */
int ofb_slab_local_rows(const ofb_slab* slab, int* first_row, int* owned_rows);
/* NCCL bootstrap: rank 0 makes the 128-byte id, the host broadcasts it (torch.distributed), every rank creates. */
int ofb_comm_unique_id(void* id128);
int ofb_comm_create(const void* id128, int rank, int world, void** comm_out);
int ofb_comm_destroy(void* comm);
/* 1 if the halo rows travel by peer stores into CUDA-IPC mailboxes (all ranks could map their neighbours),
 * 0 if the exchanges fall back to ncclSend/ncclRecv. */
int ofb_comm_peer_halo(void* comm);
size_t ofb_slab_workspace_bytes(int B, int Ny, int dtype, const ofb_slab* slab);
/* ofb_solve on a slab; every rank of slab->comm must make the same call.  Scalars, iteration counts and the
 * convergence decision are identical on all ranks. */
int ofb_solve_slab(const void* P, const void* b3, void* x3, const double* alpha, const double* beta,
                   double rtol, int max_it, const ofb_solve_options* opt, void* workspace, size_t workspace_bytes,
                   int* iters_out, double* relres_out, int* converged_out, double* recurrence_out,
                   int B, int Nx_local, int Ny, int dtype, const ofb_slab* slab, void* stream);
/* ofb_epilogue on a slab: functionals/moments are this rank's partial sums over its owned rows. */
int ofb_epilogue_slab(const void* P, const void* C, const void* x3, void* vx, void* vy, void* speed, void* gamma,
                      double* functionals, double* moments, const double* alpha, const double* beta, double scale,
                      int B, int Nx_local, int Ny, int dtype, const ofb_slab* slab, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Callers either side of the hot path (SURVEY.md section 8f NEXT-3 / NEXT-4).  Frames [F][Nx][Ny] row-major.
 * ------------------------------------------------------------------------------------------------------- */

/* Liu-Shen 2-field Jacobi iteration: replaces liu_shen_optical_flow_jit (optical_flow.py:426-673) for B frame
 *     pairs at once.  P/C = previous/current frames [B][Nx][Ny]; vx, vy: in = initial guess in pixel units
 *     (initial_v * delta_t/delta_x, :497-498), out = velocities times `scale` (= delta_x/delta_t, :669-670);
 *     speed out (:671); tmp_vx/tmp_vy = scratch of the same size.  Exactly `iterations` sweeps (the reference's
 *     `tolerance` argument is ignored there too).  Ghost values are the mirror image about the edge pixel
 *     (apply_constant_boundary_condition on the padded arrays, :1304-1316). */
int ofb_liu_shen(const void* P, const void* C, void* vx, void* vy, void* speed, void* tmp_vx, void* tmp_vy,
                 double alpha, int iterations, double scale, int B, int Nx, int Ny, int dtype, void* stream);

/* np.max over n elements (:328) -> *max_out (device double). */
int ofb_movie_max(const void* in, long long n, int dtype, double* max_out, void* stream);

/* apply_adaptive_threshold (optical_flow.py:308-338): uint8(movie/max*255), then
 *     cv2.adaptiveThreshold(.., 1.0, ADAPTIVE_THRESH_MEAN_C, THRESH_BINARY, window, threshold):
 *     out = 1 where src > boxmean(window, replicated border) - idelta, idelta = ceil(threshold).
 *     window odd, 3..255.  movie_max: device double (ofb_movie_max).  tmp_u8: scratch of 3*F*Nx*Ny bytes, 2-byte
 *     aligned (16-bit row sums of the separable box filter, then the uint8 plane). */
int ofb_adaptive_threshold(const void* in, const double* movie_max, unsigned char* tmp_u8, unsigned char* out,
                           int window, int idelta, int F, int Nx, int Ny, int dtype, void* stream);

/* apply_clahe (optical_flow.py:340-374): cv2.createCLAHE(clip_limit, (tiles_x, tiles_y)).apply on
 *     movie.astype(uint16), 65536 bins per tile; out has the dtype of in (the reference returns float64).
 *     tiles_x splits the columns, tiles_y the rows.  The workspace holds histograms and look-up tables of the
 *     frames in flight (6 bytes x 65536 x tiles per frame); the call loops over as many frames as it holds. */
size_t ofb_clahe_workspace_bytes(int frames_in_flight, int tiles_x, int tiles_y);
int ofb_clahe(const void* in, void* out, double clip_limit, int tiles_x, int tiles_y, void* workspace,
              size_t workspace_bytes, int F, int Nx, int Ny, int dtype, void* stream);

/* float32 -> float64 on the device (n elements, 16-byte aligned buffers): result planes of a float32 solve widened to
 *     the reference's float64 result format (optical_flow.py:1193-1205) before they leave for pinned host memory. */
int ofb_widen(const void* in_f32, void* out_f64, long long n, void* stream);

/* Background subtraction of conduct_optical_flow (optical_flow.py:195-198):
 *     out = movie - background where blurred > background, else 0 (blurred = ofb_blur(movie, sigma 10)). */
int ofb_subtract_background(const void* movie, const void* blurred, void* out, double background, long long n,
                            int dtype, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Sparse direct solve: replaces the reference's SuperLU branch (optical_flow.py:1146-1147, use_direct_solver=True) and
 * serves as the fallback where the operator is indefinite and no preconditioned Krylov method converges (small
 * speed_alpha).  Multifrontal LU in fp64 on a geometric nested dissection of the interior pixels; the plan (index
 * arrays that depend on the frame shape only) is built on the host (opticalflow_b200/ndplan.py) and handed in as
 * `levels`, root first: per level the padded pixel counts and DEVICE arrays of int32 pixel ids (i * (Ny-2) + j over the
 * interior, -1 = padding).  The array of structs itself lives in host memory.  Dense kernels are cuBLAS / cuSOLVER.
 * One pair is factorised and solved at a time (0.22 s per 1024^2 pair on a B200; ~28 GB of workspace).  relres_out
 * (device, B doubles, may be NULL): |b - A x| / |b| recomputed from the returned x (:1150-1151); converged_out (device,
 * B ints, may be NULL): ofb_solve's verdict on it, |b - A x| <= rtol |b| + 8 u |D x|.  x3: out only.
 * ------------------------------------------------------------------------------------------------------- */
typedef struct ofb_nd_level {
    int n_nodes, n_own, n_border, reserved;
    const int* own;             /* [n_nodes][n_own]     pixels the node eliminates */
    const int* border;          /* [n_nodes][n_border]  ring of pixels around its rectangle */
    const int* parent_slot;     /* [n_nodes][n_border]  position of each border pixel in the parent's front */
} ofb_nd_level;
size_t ofb_direct_workspace_bytes(const ofb_nd_level* levels, int n_levels, int B, int Nx, int Ny);
/* 1: the process's last direct solve replayed the CUDA graph captured for its frame shape, -1: a library call refused to be captured
 * and the solve was issued launch by launch, 0: no direct solve yet (diagnostics) */
int ofb_direct_graph_state(void);
int ofb_direct_solve(const void* P, const void* b3, void* x3, const double* alpha, const double* beta,
                     const ofb_nd_level* levels, int n_levels, void* workspace, size_t workspace_bytes,
                     double rtol, double* relres_out, int* converged_out, int B, int Nx, int Ny, int dtype, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Allocation-free variant.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * ofb_blur as the two-launch separable blur (axis 0 then axis 1, the order of scipy.ndimage.gaussian_filter), with
 * the intermediate plane in a file-scope device array of OFB_BLUR_NOALLOC_MAX_BYTES inside the library instead of a
 * caller-provided `tmp`.  The call allocates nothing (no cudaMalloc/cudaFree, no allocating library routine); the
 * caller still passes in, out and the device array of 2*radius+1 normalised weights.  Limits and sharing:
 *   - F*Nx*Ny*sizeof(dtype) must not exceed OFB_BLUR_NOALLOC_MAX_BYTES (16 MiB: four 1024^2 float32 frames, two
 *     float64 ones); larger requests return -12 with a message, use ofb_blur there;
 *   - every GPU that loads the library holds the array until the process exits;
 *   - all calls on a GPU share it and it keeps the last call's contents (the first pass overwrites every element the
 *     second reads, so nothing needs clearing); two calls that overlap on one GPU - different streams, or different
 *     host threads - race on it: serialise them. */
#define OFB_BLUR_NOALLOC_MAX_BYTES (16u << 20)
int ofb_blur_noalloc(const void* in, void* out, const double* weights, int radius,
                     int F, int Nx, int Ny, int dtype, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* OFB200_H */
