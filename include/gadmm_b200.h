/* gadmm_b200 native library -- C ABI (sm_100a).
 *
 * Every entry point returns 0 on success, a cudaError_t value (>0) or -1 (bad argument);
 * gadmm_last_error() returns the message of the calling thread's last failure.  All pointers are
 * device pointers unless stated otherwise; `stream` is a cudaStream_t passed as void*.
 * Memory is owned by the caller (PyTorch's caching allocator / symmetric memory in the Python
 * runtime); the library allocates nothing unless a function says so.
 *
 * Parity citations refer to the read-only MATLAB reference (dayoladejo/GADMM).
 */
#ifndef GADMM_B200_H
#define GADMM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GADMM_TILE_EDGE 64
#define GADMM_STORE_F64 0
#define GADMM_STORE_F32 1

const char* gadmm_last_error(void);
int gadmm_version(void);

/* ---- packed-symmetric tile layout ---------------------------------------------------------
 * A symmetric dp x dp matrix (dp = 64*nT) is stored as its lower block-triangle of 64x64
 * row-major tiles, tile (I,J), J<=I, at index I*(I+1)/2+J; diagonal tiles are stored full. */
size_t gadmm_packed_elems(int dp);

/* ---- setup (K1, K2, K12; group_ADMM_closedForm.m:39-46 `(H'*H + c*rho*I) \ ...`) ---------- */
/* G[n] = X_n' X_n (dp x dp row-major, zero padded), b[n] = X_n' y_n (dp), hyy[n] = 0.5*y_n'y_n.
 * X is [N][s][d] row-major fp64, y is [N][s]. */
int gadmm_gram_f64(const double* X, const double* y, int N, int s, int d, int dp, double* G,
                   double* b, double* hyy, void* stream);
/* A[n] = G[n] + shift*I on the leading d x d block, identity on the padded diagonal. */
int gadmm_shift_copy(const double* G, double* A, int N, int d, int dp, double shift, void* stream);
/* In-place batched SPD inverse by blocked Cholesky: A (lower block-triangle read) -> A^-1
 * (lower block-triangle written).  W: workspace N*dp*dp doubles; Dinv: N*nT*64*64 doubles. */
int gadmm_spd_inverse_f64(double* A, double* W, double* Dinv, int N, int dp, void* stream);
/* Pack the lower block-triangle of M[n] into tiles (fp64 or fp32), zeroing rows/cols >= d. */
int gadmm_pack_sym(const double* M, void* packed, int N, int d, int dp, int store, void* stream);
/* y[n] = S[n] x[n] for packed symmetric S (plain batched SYMV; used by tests, GD/DGD/LAG
 * gradients G_n theta, eig_extreme).  scratch: N*(dp + dp*dp/128) doubles (see
 * gadmm_symv_scratch_elems), counters: N uint32 zero-initialised (left zero on return). */
size_t gadmm_symv_scratch_elems(int dp);
int gadmm_symv_batched(const void* const* Sptr, const double* x, double* y, int N, int dp,
                       int store, double* scratch, unsigned int* counters, void* stream);

/* Same product with no working memory supplied by the caller (no counterpart in the reference;
 * added for callers that cannot allocate device memory: plug-ins inside an application that owns
 * the whole device arena, or latency-critical loops that must avoid cudaMalloc/cudaFree).  The
 * partial sums and ticket counters live in a fixed array inside the library (held on every GPU
 * that loads it until the process exits).  Limits: N <= GADMM_NOALLOC_MAX_MATRICES and
 * N*gadmm_symv_scratch_elems(dp) <= GADMM_NOALLOC_SCRATCH_DOUBLES; beyond them the call fails
 * with a message and allocates nothing.  All calls on a device share the array: calls that can
 * overlap on the GPU race -- keep them on one stream.  The first call on a device performs the
 * one-time set-up and should be made before any stream capture. */
#define GADMM_NOALLOC_MAX_MATRICES 4096
#define GADMM_NOALLOC_SCRATCH_DOUBLES (1u << 20)
size_t gadmm_symv_noalloc_max_scratch_elems(void);
int gadmm_symv_noalloc_max_matrices(void);
int gadmm_symv_batched_noalloc(const void* const* Sptr, const double* x, double* y, int N, int dp,
                               int store, void* stream);

/* ---- chain GADMM / D-GADMM hot loop (K3, K4, K5) ------------------------------------------
 * One half-iteration = one launch of the fused kernel: RHS assembly
 * r = b_n - nu_n + rho*(theta_l + theta_r), theta_n = M_n r by a packed-symmetric SYMV streamed
 * with cp.async.bulk + mbarrier, per-worker objective, store of theta_n into the local replica
 * and -- for workers whose neighbours live on other GPUs -- straight into the peers' replicas
 * over NVLink followed by a release flag.  (group_ADMM_closedForm.m:13-57 head pass, :61-92 tail
 * pass; node-dual form static_group_ADMM_closedForm.m:50-80.)  The dual kernel applies
 * nu_n += rho*(theta_n-theta_r) - rho*(theta_l-theta_n) (static_group_ADMM_closedForm.m:130-144),
 * reduces the objective across GPUs through peer stores, evaluates the stop test
 * (group_ADMM_closedForm.m:96-108) and advances the device-resident iteration counter. */
typedef struct GadmmChain {
  /* per local worker, length Nloc */
  const int* gid;            /* global worker id */
  const int* left;           /* global id of chain predecessor, -1 at the chain start */
  const int* right;          /* global id of chain successor, -1 at the chain end */
  const int* color;          /* 0 = head group, 1 = tail group */
  const int* cmul;           /* c_n: 1 at chain ends, 2 inside */
  const void* const* Mptr;   /* packed (G_n + c_n*rho*I)^-1 to use this iteration */
  const int* push0;          /* peer rank that must receive theta_n (or -1) */
  const int* push1;
  const int* owner;          /* [Nglob] owning rank of every worker */
  const double* b;           /* [Nloc][dp]  X_n' y_n */
  const double* hyy;         /* [Nloc]      0.5*y_n'y_n */
  double* nu;                /* [Nloc][dp]  node duals */
  double* theta;             /* [2][Nglob][dp] local replica, slot = version & 1 */
  unsigned long long* flags; /* [Nglob] version of theta_g present in the local replica */
  double* const* theta_peer; /* [nranks] replica base pointer on every rank (self included) */
  unsigned long long* const* flags_peer; /* [nranks] */
  double* objpart;           /* [4][nranks] per-rank objective partials (ring over iterations) */
  unsigned long long* objflag; /* [4][nranks] */
  double* const* objpart_peer;            /* [nranks] */
  unsigned long long* const* objflag_peer; /* [nranks] */
  double* rowpart;           /* scratch [Nloc][nsplit][dp] */
  double* colpart;           /* scratch [Nloc][dp*dp/128] */
  unsigned int* counters;    /* [Nloc+2] zero-initialised */
  double* objw;              /* [2][Nloc] per-worker objective, slot = iteration & 1 */
  unsigned long long* ctrl;  /* [4]: ctrl[0]=iteration about to run (1-based), ctrl[1]=Iter at
                                which loss<acc first held (0 = not yet), ctrl[2]=stop enabled */
  double* obj_hist;          /* [max_iter+1] objective per iteration (index = iteration) */
  double* loss_hist;         /* [max_iter+1] */
  unsigned long long* t_hist; /* [max_iter+1] %globaltimer at the end of each iteration */
  double rho;
  double obj0;
  double acc;
  int d, dp, Nglob, Nloc, rank, nranks, store, max_iter;
  int pdl;                   /* 1: every launch of a window after its first head phase uses
                                programmatic dependent launch; all ordering on one GPU then goes
                                through the version flags, like the ordering between GPUs */
  int split_finish;          /* 1: the column-partial reduction / theta publication of every
                                worker runs in gadmm_finish_kernel (one 64-thread CTA per 64
                                entries) instead of in the worker's last streaming CTA */
  unsigned long long* wready; /* [Nloc] iteration whose partial sums are complete (split_finish) */
  double* sblk;              /* [Nloc][dp/64][2] per-block objective pieces (split_finish) */
  unsigned int* fin_counters; /* [Nloc] zero-initialised (split_finish) */
  int nsplit;                /* >= 1: CTAs per pair of tile-rows (tile columns divided into nsplit
                                ranges); rowpart is then [Nloc][nsplit][dp] */
  int reserved;
  const int* order;          /* [Nloc] permutation of the local workers: launch order of their CTAs
                                (workers with a neighbour on another GPU first) */
  unsigned long long* sync;  /* [2]: [0] = first iteration of the window being executed, [1] = last
                                iteration whose stop decision is final (both written on the device) */
  unsigned long long* nuflag; /* [Nloc] iteration through which nu_k has been updated */
} GadmmChain;

int gadmm_chain_half_iter(const GadmmChain* c, int color, void* stream);
int gadmm_chain_dual_objective(const GadmmChain* c, void* stream);
/* Convenience: n_iter x (head, tail, dual) on `stream` (capturable into a CUDA graph). */
int gadmm_chain_iterate(const GadmmChain* c, int n_iter, void* stream);
/* Set the kernels' shared-memory attributes up front (call once before a CUDA-graph capture). */
int gadmm_chain_configure(void);
/* Number of kernel launches gadmm_chain_iterate issues per iteration. */
int gadmm_chain_launches_per_iter(void);

/* Persistent single-launch variant for small problems (one GPU, fp64 tiles, dp <= 128, one CTA
 * per worker co-resident; cooperative launch).  The whole state lives in shared memory, phase
 * ordering is by neighbour flags, a grid barrier every 8 iterations agrees on the stop decision.
 * objring: 32*Nloc doubles, ringcnt: 32 uint32 (zero-initialised); ctrl[3] = ~0 before first use. */
int gadmm_chain_persistent_supported(const GadmmChain* c);
int gadmm_chain_persistent(const GadmmChain* c, int n_iter, double* objring, unsigned int* ringcnt,
                           void* stream);

/* ---- D-GADMM re-chaining on the device (K10; dynamic_group_ADMM_closedForm.m:18-21) -----------
 * Launched once per iteration in front of the head kernel; a no-op unless
 * it > 1 && it % tau == 0.  Window index kk = it/tau - (tau==1) - dyn[0] selects paths[kk]. */
typedef struct GadmmRechain {
  const int* paths;          /* [K][N] chains (0-based worker ids in chain order) */
  int* pos;                  /* [N] scratch */
  int* tab;                  /* [7][Nloc] gid,left,right,color,cmul,push0,push1 (GadmmChain tables) */
  long long* Mptr;           /* [Nloc] */
  const int* owner;          /* [N] */
  double* theta;             /* local replica [2][N][dp] */
  unsigned long long* flags; /* [N] */
  double* const* theta_peer; /* [nranks] */
  const unsigned long long* ctrl;
  long long base_c1, base_c2, stride_bytes;   /* packed inverses for c=1 / c=2, bytes per worker */
  const long long* dyn;      /* device [2]: window index of paths[0], number of valid rows */
  long long tau;
  int N, Nloc, dp, rank, nranks, K;
  int* order;                /* optional [Nloc]: GadmmChain.order, rewritten to chain order on a re-chain */
} GadmmRechain;
int gadmm_rechain(const GadmmRechain* p, void* stream);

/* ---- first-order baselines (K7, K8; GD_DGD_LAG.m, GD_DGD_LAG_logistic.m, dual_averaging*.m) ---- */
/* Gradient providers.  Linear: Y[n] = S_n x_n - b_n, fw[n] = 0.5 x'S x - b'x + hyy[n]; x_stride
 * = 0 shares one iterate between all workers (parameter-server methods). */
int gadmm_symv_grad(const void* const* Sptr, const double* x, long long x_stride, const double* b,
                    const double* hyy, double* Y, double* fw, int N, int dp, int store,
                    double* scratch, unsigned int* counters, void* stream);
/* Logistic: Y[n] = -X_n'(y./(1+exp(y.*X_n th))) + lam*th, fw[n] = lam/2||th||^2 + sum log(1+exp(-y.*z)),
 * wout (optional) = sigma(1-sigma) per sample; wbuf: N*s doubles of scratch. */
int gadmm_logistic_grad(const double* X, const double* y, const double* th, long long th_stride,
                        double lam, int N, int s, int d, int dp, double* Y, double* fw,
                        double* wout, double* wbuf, void* stream);

#define GADMM_FO_GD 0
#define GADMM_FO_LAG_PS 1
#define GADMM_FO_LAG_WK 2
#define GADMM_FO_IAG_CYCLIC 3
#define GADMM_FO_IAG_RANDOM 4

typedef struct GadmmShared {      /* shared-iterate methods: GD, LAG-PS, LAG-WK, IAG */
  double* theta;                  /* [dp] */
  const double* Y;                /* [N][dp] fresh gradients at theta */
  const double* fw;               /* [N] per-worker objective at theta */
  double* ghat;                   /* [N][dp] server-side (possibly stale) gradients, init ones */
  double* theta_hat;              /* [N][dp] iterate of each worker's last upload (LAG-PS) */
  double* dnorm;                  /* [10] ring of ||theta^{k+1}-theta^k||^2 */
  const double* Hmax;             /* [N] */
  const double* uni;              /* [max_iter+2] uniforms for R-IAG (index = iteration) */
  const double* cumprob;          /* [N] cumulative sampling probabilities */
  double* obj_hist;               /* [max_iter+2] */
  double* loss_hist;
  double* comm_hist;              /* comm_iter_final */
  unsigned long long* t_hist;
  unsigned long long* ctrl;       /* [0]=iteration (1-based), [1]=hit, [2]=stop enabled, [3]=comm_iter */
  unsigned int* counters;         /* [2] ticket, uploads of this iteration (zero-initialised) */
  /* multi-GPU / parallel server (split_server != 0): workers gid0..gid0+N-1 of Nglob live here;
   * column sums, objective partial and upload count are exchanged through xbuf/xflag (symmetric
   * memory, P2P stores + release flags); the theta step is replicated on every GPU. */
  double* xbuf;                   /* [2][nranks][dp+64] */
  unsigned long long* xflag;      /* [2][nranks] */
  double* const* xbuf_peer;       /* [nranks] */
  unsigned long long* const* xflag_peer;
  double* dnpart;                 /* [dp/64] */
  unsigned int* counters2;        /* [2] zero-initialised */
  double step, thr, obj0, acc;
  int N, d, dp, kind, max_iter, logistic;
  int Nglob, gid0, rank, nranks, split_server, reserved;
} GadmmShared;
int gadmm_shared_update(const GadmmShared* p, void* stream);

typedef struct GadmmOwn {         /* one iterate per worker: DGD, dual averaging */
  double* Theta;                  /* [N][dp] */
  const double* Y;                /* [N][dp] fresh gradients at Theta */
  const double* fw;               /* [N] */
  double* Z;                      /* [N][dp] dual-averaging accumulators (Z_prev) */
  double* obj_hist;
  double* loss_hist;
  unsigned long long* t_hist;
  unsigned long long* ctrl;       /* [0]=iteration, [1]=hit, [2]=stop enabled */
  unsigned int* counters;         /* [1] */
  /* multi-GPU (nranks > 1): workers gid0..gid0+N-1 of Nglob live here (contiguous ranges).  The
   * tables below are symmetric allocations indexed by GLOBAL worker id; a GPU writes its own rows
   * locally and pushes its first / last row into the neighbouring GPU's table (NVLink P2P stores
   * + st.release.sys row flags) -- the chain halo of SURVEY.md K8. */
  double* Ytab;                   /* [Nglob][dp] gradients; Y == Ytab + gid0*dp */
  double* Ztab;                   /* [2][Nglob][dp] halo copies of Z, slot = version & 1 */
  unsigned long long* rowflag;    /* [2][Nglob]: version of the Y / Z halo row present here */
  double* const* Ytab_peer;       /* [nranks] */
  double* const* Ztab_peer;
  unsigned long long* const* rowflag_peer;
  double* fx;                     /* [2][nranks] objective partials (ring over iterations) */
  unsigned long long* fxflag;     /* [2][nranks] */
  double* const* fx_peer;
  unsigned long long* const* fxflag_peer;
  double* ftot;                   /* [2] device scalars: all-rank objective of iteration it */
  double step, obj0, acc;
  int N, d, dp, max_iter, jacobi;
  int Nglob, gid0, rank, nranks, reserved;
} GadmmOwn;
/* Objective exchange (+ push of the boundary gradient rows when push_y != 0); a no-op for
 * nranks == 1.  Launch after the gradient provider, before the update kernel. */
int gadmm_own_exchange(const GadmmOwn* p, int push_y, void* stream);
int gadmm_dgd_update(const GadmmOwn* p, void* stream);
/* update=1: record the objective of the previous iterate (from fw), stop test, then one dual
 * averaging sweep; update=0: record only (closes the last iteration). */
int gadmm_dualavg_update(const GadmmOwn* p, int update, void* stream);

/* ---- standard (star) ADMM (K6; standared_ADMM.m:17-107) ------------------------------------
 * Four launches per iteration (workers, reduce, server stream, z finisher).  Workers kernel: theta_n = M_n (b_n - lambda_n + rho z) for the
 * N-1 workers (lambda_n += rho(theta_n - z) of the previous iteration is folded into the RHS
 * assembly), per-worker objective, and -- by the last CTA -- the reduction
 * u = sum_n (lambda_n + rho theta_n), pushed with P2P stores + a release flag into every GPU.
 * Server kernel (replicated on every GPU, so no broadcast is needed): waits for all ranks'
 * partials, z = M_s (b_s + u), total objective, stop test, iteration counter. */
typedef struct GadmmStar {
  const void* const* Mptr;   /* [Nw] packed (G_n + rho I)^-1 of the local workers */
  const void* Ms;            /* packed (G_s + (N-1) rho I)^-1 of the server shard (replicated) */
  const double* b;           /* [Nw][dp] */
  const double* hyy;         /* [Nw] */
  const double* bs;          /* [dp] server shard */
  double* lam;               /* [Nw][dp] */
  double* theta;             /* [Nw][dp] */
  double* z;                 /* [dp] */
  double* U;                 /* [Nw][dp] scratch: lambda_n + rho theta_n */
  double* fw;                /* [Nw] */
  double* ubuf;              /* symmetric: [2][nranks][dp+64] partial u (+ objective at [dp]) */
  unsigned long long* uflag; /* symmetric: [2][nranks] */
  double* const* ubuf_peer;  /* [nranks] */
  unsigned long long* const* uflag_peer;
  double* rowpart;           /* [Nw+1][dp] (last slot: server solve) */
  double* colpart;           /* [Nw+1][colpart elems] */
  unsigned int* counters;    /* [Nw+2] zero-initialised */
  unsigned long long* ctrl;  /* [0]=iteration, [1]=hit, [2]=stop enabled */
  double* obj_hist;
  double* loss_hist;
  unsigned long long* t_hist;
  /* NVLS mode (nvls != 0): partial sums stay local and are reduced in the NVSwitch with
   * multimem.ld_reduce by the server GPU only, which multicasts z and the objective back
   * (multimem.st) and raises zflag on every other GPU. */
  const double* ubuf_mc;     /* multicast alias of ubuf */
  double* z_mc;              /* multicast alias of z (z must then live in symmetric memory) */
  double* zobj;              /* symmetric: [4] objective broadcast ring */
  double* zobj_mc;
  unsigned long long* zflag; /* symmetric: [1] */
  unsigned long long* const* zflag_peer; /* [nranks] */
  double rho, obj0, acc, hyy_s;
  int d, dp, N, Nw, rank, nranks, store, max_iter, nvls, server_rank;
  int nsplit;                /* CTAs per tile-row pair of the server solve (0/1 = one); rowpart then
                                holds [Nw + nsplit][dp] and counters [Nw + 4] */
  int reserved;
  double* zblk;              /* [dp/64][2] per-block objective pieces of the server solve */
} GadmmStar;
int gadmm_star_iterate(const GadmmStar* p, int n_iter, void* stream);

/* ---- logistic GADMM, inexact local solves (K9; group_ADMM_logistic_GD.m, logReg_GD.m) ------- */
typedef struct GadmmLogistic {
  const double* X;           /* [nloc][s][d] */
  const double* y;           /* [nloc][s]   labels +-1 */
  double* theta;             /* [N][dp] */
  double* lam;               /* [N][dp]  edge duals: lam[n] couples workers n and n+1 */
  const double* fw;          /* [N] per-worker objective (from gadmm_logistic_grad) */
  unsigned long long* inner_steps; /* optional: total inner GD steps */
  unsigned long long* ctrl;  /* [0]=iteration, [1]=hit, [2]=stop enabled */
  unsigned int* counters;    /* [1] zero-initialised */
  double* obj_hist;
  double* loss_hist;
  unsigned long long* t_hist;
  double rho, lam_log, gd_step, inner_tol, obj0, acc;
  int N, s, d, dp, max_iter, max_inner;
  int n0, nloc;              /* half_gd handles workers n0..n0+nloc-1 (nloc <= 0: all N); X, y, fw hold
                                only those shards, theta / lam are [N][dp] tables (multi-GPU: replicas) */
} GadmmLogistic;
int gadmm_logistic_half_gd(const GadmmLogistic* p, int color, void* stream);
int gadmm_logistic_dual_obj(const GadmmLogistic* p, void* stream);

/* Hessian-free logistic building block for shards too wide for an explicit Hessian (s < d, d up to
 * 8192): v[n] = X_n' w(X_n x_n) in ONE pass over the fp32 shard X: [N][s][d].
 *   hv = 0: w = -y ./ (1 + exp(y .* Xx)); also writes Dw = sigma(1-sigma) [N][s] and loss[n] =
 *           sum log(1 + exp(-y .* Xx))   (data part of the gradient, group_ADMM_logistic.m:48-55)
 *   hv = 1: w = Dw .* (X x)              (v = X' D X x, the Hessian product of Newton-CG)
 * x has row stride x_stride doubles (0 = one vector shared by all shards), v row stride ldv.
 * Shard n of the call is shard n*shard_stride of X / y / Dw (2 = one colour of a chain).
 * G = CTAs per shard; scratch = gadmm_logistic_xtwx_scratch_elems(N, d, G) doubles. */
size_t gadmm_logistic_xtwx_scratch_elems(int N, int d, int G);
int gadmm_logistic_xtwx(const float* X, const double* y, const double* x, long long x_stride,
                        double* Dw, double* v, int ldv, double* loss, int N, int s, int d, int G,
                        int hv, int shard_stride, double* scratch, void* stream);

/* Same product with fp64 shards (`store` = GADMM_STORE_F64) or fp32 shards (GADMM_STORE_F32): user
 * data that is not exactly representable in fp32 stays in double, as the reference computes
 * (group_ADMM_logistic.m:48-49).  hv = 2 (fp32 multiplies) applies to fp32 shards only. */
int gadmm_logistic_xtwx_ex(const void* X, int store, const double* y, const double* x,
                           long long x_stride, double* Dw, double* v, int ldv, double* loss, int N,
                           int s, int d, int G, int hv, int shard_stride, double* scratch,
                           void* stream);

/* ---- logistic GADMM with exact local solves, device-resident (K9; group_ADMM_logistic.m:48-55
 * head solves, :83-86 tail solves, :90-91 dual update, :96-110 objective / stop) ----------------
 * Every worker's strongly convex sub-problem is solved by a batched Newton-CG whose vector
 * recurrences, forcing-term test and Armijo line search run in kernels around the one-pass
 * products above.  The whole run is ONE CUDA graph launch: WHILE nodes (conditional graph nodes,
 * their conditions set on the device) express iteration / Newton / CG / line-search loops, so
 * there is no host round trip, no torch op and no NCCL call between the first iteration and the
 * stop.  Boundary models are stored into the neighbouring GPUs' tables over NVLink followed by a
 * release flag; the dual update and the all-rank objective sum use the same flags. */
typedef struct GadmmLogi {
  const void* X;             /* [Nloc][s][d] shards, fp32 or fp64 (`store`) */
  const double* y;           /* [Nloc][s] labels +-1 */
  double* Dw;                /* [Nloc][s] sigma(1-sigma) at the last evaluated point */
  double* table;             /* symmetric: [N][d] model of every worker (own rows + halo rows) */
  unsigned long long* flags; /* symmetric: [N] iteration of the model present in `table` */
  double* const* table_peer; /* [nranks] */
  unsigned long long* const* flags_peer;
  double* objpart;           /* symmetric: [4][nranks] objective partials (ring over iterations) */
  unsigned long long* objflag;
  double* const* objpart_peer;
  unsigned long long* const* objflag_peer;
  double* lam;               /* [Nloc+1][d] edge duals: lam[k] couples global workers g0-1+k, g0+k */
  /* work vectors of one colour's batched solve, rows = that colour's local workers: [R][d] each */
  double* x;
  double* g;
  double* p;
  double* r;
  double* q;
  double* hq;                /* Hessian product / trial gradient data part (output of the pass) */
  double* xtry;
  double* lin;
  double* vlast;             /* [2][R][d] data part of the gradient at each colour's current models */
  /* per-row scalars [R] (losslast: [2][R]) */
  double* mu;
  double* f;
  double* gn;
  double* eta;
  double* rs;
  double* t;
  double* slope;
  double* losstry;
  double* losslast;
  int* live;
  double* objw;              /* [Nloc] */
  double* scratch;           /* partial column sums of the pass: xtwx scratch for max(R0,R1) rows */
  unsigned long long* ctrl;  /* [16]: 0 iteration about to run, 1 Iter at which loss<acc first held,
                                2 stop enabled, 3 last iteration of this launch, 4 newton count,
                                5 cg count, 6 line-search count, 8 total Hessian products,
                                9 total evaluations, 10 total Newton steps, 11 / 12 / 13 ns spent
                                waiting for neighbour models (solves / dual step) and for the
                                other ranks' objective partials */
  unsigned int* cond;        /* [4] loop conditions as last written (outer, newton, cg, ls) */
  unsigned int* counters;    /* [8] zero-initialised tickets */
  double* obj_hist;          /* [max_iter+2] */
  double* loss_hist;
  unsigned long long* t_hist;
  double rho, lam_log, obj0, acc, tol;
  int N, Nloc, g0, s, d, rank, nranks, store, hv_mode;
  int max_iter, max_newton, max_cg, max_ls;
  int G[2];                  /* CTAs per shard of the streaming pass, per colour */
} GadmmLogi;
/* 0 when conditional graph nodes work on this driver / device (probe: a WHILE loop counting to n). */
int gadmm_graph_loops_selftest(int n, void* stream);
/* Rows of colour c on this rank (R = max over the two colours sizes the work vectors). */
int gadmm_logi_rows(const GadmmLogi* p, int colour);
/* Evaluate both colours at the models in `table` (fills vlast / losslast / Dw); call once before
 * the first run and after the models were changed from outside. */
int gadmm_logi_prime(const GadmmLogi* p, void* stream);
/* Build the graph for `p` (pointers and scalars are baked in).  Returns an opaque handle. */
int gadmm_logi_graph_create(const GadmmLogi* p, void** handle_out);
int gadmm_logi_graph_launch(void* handle, void* stream);
int gadmm_logi_graph_destroy(void* handle);
/* Same iteration driven from the host (one condition read-back per loop test); for debugging and
 * as the fall-back where conditional nodes are unavailable.  Runs until ctrl[3] or the stop. */
int gadmm_logi_run_host(const GadmmLogi* p, void* stream);

/* ---- NVLink flag-latency probes (no counterpart in the reference; profiles/flag_latency.md) ----
 * pingpong: role 0 stores a word + release flag into the peer and waits for the answer, role 1
 * answers; out_ns[0] = device time of `iters` round trips.  exchange: the all-rank scalar exchange
 * of the dual/objective kernels (lane q <-> rank q), `iters` times; ring buffers as GadmmChain's
 * objpart/objflag ([4][nranks]); `base` = last iteration number already used on these flags. */
int gadmm_probe_pingpong(unsigned long long* my_flag, double* my_word, unsigned long long* peer_flag,
                         double* peer_word, int iters, int role, unsigned long long* out_ns,
                         void* stream);
int gadmm_probe_exchange(double* part, unsigned long long* flag, double* const* part_peer,
                         unsigned long long* const* flag_peer, int rank, int nranks, int iters,
                         unsigned long long base, unsigned long long* out_ns, double* out_sum,
                         void* stream);

/* ---- tensor-core Gram (K1 proper): G[n] (+)= X_n' diag(w_n) X_n with tcgen05 kind::tf32 in a
 * 3-product hi/lo split (fp32-grade), TMA-fed, TMEM accumulators, fp64 result; K-segments are summed
 * in 64-bit fixed point (integer atomics), so the result is bit-reproducible.  X: [N][s][d] fp64
 * or fp32; w: optional [N][s] row weights (logistic Hessian X'DX); work: gadmm_gram_tc_workspace_floats
 * floats; kseg: samples per fp32 accumulation segment (0 = 2048). */
size_t gadmm_gram_tc_workspace_floats(int N, int s, int d);
int gadmm_gram_tcgen05(const void* X, int x_is_f32, const double* w, int N, int s, int d, int dp,
                       float* work, double* G, int accumulate, int kseg, void* stream);

/* ---- topology and synthetic data on the device (K10, K13, K14) -------------------------------
 * Philox4x32-10 fills (seed, stream_id) -> reproducible; out[i] = scale*U[0,1) / N(mean, sd). */
int gadmm_philox_uniform(double* out, size_t n, unsigned long long seed, unsigned int stream_id,
                         double scale, void* stream);
int gadmm_philox_normal(void* out, size_t n, int is_f32, unsigned long long seed,
                        unsigned int stream_id, double mean, double sd, void* stream);
/* Greedy nearest-neighbour chain from node 0 over positions xy[N][2] (findPath.m:16-30; model 1 =
 * findPath2.m: energy costs, P_central, centre node).  path: N ints (0-based), cost: N-1,
 * d_square (optional): N*N, P_central (optional): N, center (optional): 1 int. */
int gadmm_find_path(const double* xy, int N, int model, int* path, double* cost, double* d_square,
                    double* P_central, int* center, void* stream);
/* K chains at once (one CTA each): xy [K][N][2] -> path [K][N], cost [K][max(N-1,1)]. */
int gadmm_find_path_batch(const double* xy, int K, int N, int model, int* path, double* cost,
                          void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GADMM_B200_H */
