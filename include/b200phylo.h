/*
 * b200phylo.h — C ABI of libb200phylo.so: tree log-likelihood + branch-length gradient
 * on one NVIDIA B200 (sm_100a).  Plain pointers and sizes only; no C++/torch types.
 *
 * What this boundary replaces.  /root/reference holds no likelihood code: its
 * contenders reach a native engine through (i) BEAGLE's C API as driven by bito /
 * torchtree-bito (selected with `--engine bitorch`, /root/reference/modules/macro_flu.nf:119-122,
 * and by `bitorch-benchmark`, /root/reference/modules/micro.nf:8,47-53) and (ii)
 * physher's treelikelihood (`physher-benchmark`, /root/reference/modules/micro.nf:20-24).
 * Each entry point below names the BEAGLE-style call an integrator would otherwise
 * bind (upstream recall, not verifiable on this box — SURVEY.md §2.1) and the
 * reference JSON / CLI element that feeds it.
 *
 * Conventions
 *   - every function returns B2P_OK (0) or a negative error code; never throws;
 *     b2p_last_error() returns the message for the last failure on that instance
 *     (pass NULL for failures of b2p_create itself).
 *   - node ids: tips 0..N-1, internal nodes N..2N-2 in post-order, root = 2N-2;
 *     a branch is named by its child node id, so branch arrays have length 2N-2.
 *   - pointers called h_* are host memory, d_* are device memory on the instance's
 *     device.  The instance owns all scratch (partials, matrices, schedules); the
 *     caller owns every buffer it passes in.  Setup calls (b2p_set_*) copy from host
 *     and synchronise; evaluation calls with a stream are asynchronous on that stream
 *     and perform no hidden synchronisation.
 *   - one instance = one device; an instance is not thread-safe; instances are
 *     independent.  There is no CPU fallback: creation fails without a CUDA device.
 */
#ifndef B200PHYLO_H
#define B200PHYLO_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2P_OK 0
#define B2P_ERR_ARG (-1)       /* bad argument / call order */
#define B2P_ERR_CUDA (-2)      /* CUDA runtime error (message has the cudaError string) */
#define B2P_ERR_ALLOC (-3)     /* host or device allocation failed */
#define B2P_ERR_UNSUPPORTED (-4)

#define B2P_PRECISION_F64 64
#define B2P_PRECISION_F32 32

typedef struct b2p_instance b2p_instance;

typedef struct b2p_config {
  int32_t n_tips;       /* N >= 2 */
  int32_t n_patterns;   /* P >= 1 (this rank's slice when sharded) */
  int32_t n_categories; /* K >= 1 rate categories, K <= 32 */
  int32_t n_states;     /* 4 (nucleotide walk kernel) or 5..64 (dense-state path, e.g. 61 codons:
                           one rate category, fp64) */
  int32_t precision;    /* B2P_PRECISION_F64 | B2P_PRECISION_F32 (partials / matrices) */
  int32_t device;       /* CUDA device ordinal */
  int32_t warps_per_block; /* 0 = auto; threads per CTA / 32 (32, 24, 16 or 8) */
  int32_t flags;        /* 0, or B2P_FLAG_OUTER_PRODUCTS */
} b2p_config;

typedef struct b2p_info {
  int32_t n_tips, n_patterns, n_categories, n_states, precision;
  int32_t warps_per_block;  /* threads per CTA / 32 */
  int32_t cols_per_tile;    /* (pattern,category) columns one CTA walks the tree for at a time */
  int32_t n_tiles;          /* column tiles = independent traversals */
  int32_t grid_blocks;      /* persistent CTAs launched */
  int32_t blocks_per_sm;    /* residency the launch was sized for */
  int32_t n_sms;
  int32_t post_rounds;      /* schedule length of the post-order pass */
  int32_t tree_depth;       /* longest root-to-cherry path, in internal nodes */
  int32_t regs_per_thread;  /* of the traversal kernel in use */
  int32_t tasks_per_round;  /* node tasks a CTA executes side by side */
  int32_t pre_rounds;       /* schedule length of the pre-order pass */
  int32_t pool_blocks;      /* shared-memory blocks the planner allocates operands / results in */
  int32_t smem_bytes;       /* dynamic shared memory per CTA */
  int32_t copy_depth;       /* rounds a bulk load is issued ahead of its use */
  int32_t reserved0;
  int64_t device_bytes;     /* scratch owned by the instance */
  int64_t alg_bytes_loglik; /* (2N-3)U + N*P */
  int64_t alg_bytes_grad;   /* (5N-9)U + 2*N*P */
  int64_t resident_reads;   /* per tile: operand blocks read where they were produced (no copy) */
  int64_t block_loads;      /* per tile: operand blocks bulk-loaded from HBM/L2 */
  int64_t block_stores;     /* per tile: result blocks bulk-stored to HBM */
} b2p_info;

/* BEAGLE: beagleCreateInstance / beagleFinalizeInstance. */
int b2p_create(const b2p_config* cfg, b2p_instance** out);
int b2p_destroy(b2p_instance* h);
const char* b2p_last_error(const b2p_instance* h);
int b2p_version(void);
int b2p_get_info(const b2p_instance* h, b2p_info* out);

/* Tree topology: h_triples int32 [N-1][3] rows (node, left, right) in post-order.
 * Fed by "tree_model" (/root/reference/flu_H3N2/phylotorch-JC69.template:14-44) and
 * `-t <newick>` (/root/reference/modules/micro.nf:20,47).  BEAGLE: the operations
 * array of beagleUpdatePartials / beagleUpdatePrePartials.  Builds the static
 * schedule of both passes (list scheduling, Hu's level rule), the shared-memory
 * block allocation and the per-round copy lists of the traversal kernel. */
int b2p_set_topology(b2p_instance* h, const int32_t* h_triples);

/* Tip data: h_masks uint8 [N][P], bit i set = state i compatible (A=1,C=2,G=4,T=8;
 * gap/N = 15).  For n_states > 4 the bytes are state CODES (0..n_states-1; any other
 * value = missing data).  Fed by "site_pattern" (/root/reference/flu_H3N2/phylotorch-JC69.template:59-69,
 * /root/reference/flu_H3N2/physher-JC69.template:11-21).  BEAGLE: beagleSetTipStates /
 * beagleSetTipPartials.
 * There is deliberately no b2p_set_tip_partials: every tip partial on the reference's path is a
 * 0/1 vector built from a sequence character — torchtree's SitePattern does exactly that, and
 * physher with "tipstates": false (/root/reference/flu_H3N2/physher-JC69.template:11) stores the
 * same 0/1 vectors as partials instead of state indices — and a 0/1 vector over 4 states IS a
 * mask (over more states: a code or "missing").  Real-valued tip partials (sequencing-error or
 * uncertainty models) appear nowhere in the benchmark; they would need a float tip-row layout in
 * the traversal kernel, which the hot path should not pay for. */
int b2p_set_tip_states(b2p_instance* h, const uint8_t* h_masks);

/* Pattern weights: h_weights double [P].  BEAGLE: beagleSetPatternWeights. */
int b2p_set_pattern_weights(b2p_instance* h, const double* h_weights);

/* Substitution model as an eigen system Q = V diag(lambda) V^-1, row-major [S][S],
 * plus stationary frequencies [S] (S = n_states).  Fed by "substitution_model"
 * (/root/reference/flu_H3N2/phylotorch-JC69.template:70-73) and `--gtr`
 * (/root/reference/modules/micro.nf:52).  BEAGLE: beagleSetEigenDecomposition +
 * beagleSetStateFrequencies. */
int b2p_set_eigen(b2p_instance* h, const double* h_evec, const double* h_ivec,
                  const double* h_evals, const double* h_freqs);

/* Rate categories: h_rates, h_weights double [K].  Fed by "site_model"
 * (/root/reference/flu_H3N2/phylotorch-JC69.template:45-48).  BEAGLE:
 * beagleSetCategoryRates + beagleSetCategoryWeights. */
int b2p_set_categories(b2p_instance* h, const double* h_rates, const double* h_weights);

/* Log-likelihood only.  d_bl: device double [2N-2] branch lengths (substitutions per
 * site, by child node id); d_logl: device double [1].  BEAGLE:
 * beagleUpdateTransitionMatrices + beagleUpdatePartials + beagleCalculateRootLogLikelihoods.
 * Times the "evaluation" rows of the micro CSV (/root/reference/plot.Rmd:353). */
int b2p_loglik(b2p_instance* h, const double* d_bl, double* d_logl, void* cuda_stream);

/* Log-likelihood and d logL / d branch length for all 2N-2 branches.  d_grad: device
 * double [2N-2].  BEAGLE: the above + beagleUpdatePrePartials +
 * beagleCalculateEdgeDerivatives (bito phylo_gradients).  Times the "gradient" rows
 * (/root/reference/plot.Rmd:362). */
int b2p_loglik_grad(b2p_instance* h, const double* d_bl, double* d_logl, double* d_grad,
                    void* cuda_stream);

/* Host-buffer form of b2p_loglik_grad: copies h_bl in, runs, copies logL and the
 * gradient out and synchronises.  h_grad may be NULL (log-likelihood only).  This is
 * the call a ctypes / cgo / pybind integrator with CPU-resident parameters makes. */
int b2p_loglik_grad_host(b2p_instance* h, const double* h_bl, double* h_logl,
                         double* h_grad);

/* Gradients w.r.t. substitution-model parameters come out of the same
 * two passes.  The pre-order pass contracts, per branch b and category k, the pre-order
 * vector q with A m (m = P x, the branch's post-order message); for branch lengths A = Q and
 * the category scale is its rate.  With another operator the "gradient" output becomes
 *     g_b = sum_p (w_p / L_p) sum_k c_k s_k  q_{b,k}^T A m_{b,k}          (c_k category weights)
 * — with A = E_ij (a single 1) and s = a one-hot vector that is entry (i,j) of the outer-
 * product sum G_{b,k} = sum_p (w_p / L_p) c_k q m^T, and 16 K evaluations give every G_{b,k}.
 * Any parameter theta of Q, of the frequencies or of the category rates then has
 *     dlogL/dtheta = sum_{b,k} < G_{b,k} P_{b,k}^{-T} , dP_{b,k}/dtheta > + sum_i (dpi_i/dtheta) G_{root child}[i][i] / pi_i,
 * a few thousand 4x4 products the caller does (b200phylo/paramgrad.py).  h_op: 16 doubles,
 * row-major (NULL restores Q and the rates); h_scale: K doubles (NULL = all 1). */
int b2p_set_gradient_operator(b2p_instance* h, const double* h_op, const double* h_scale);

/* All of G in ONE evaluation.  An instance created with B2P_FLAG_OUTER_PRODUCTS in
 * b2p_config.flags (4 states, fp64, 1 or 4 rate categories) runs the walk with one column per
 * thread; b2p_branch_outer_products then also sums, in the pre-order pass, the two 4x4 outer
 * products of every task over its columns — per warp on the fp64 tensor cores (the columns are
 * the contraction index of an m8n8k4 product), per task by adding the warps' shared-memory rows
 * at the top of the next round, per evaluation by a reduction over CTAs — and writes
 *     d_outer [K][2N-2][4][4]:  G_{b,k}[i][j] = sum_p (w_p / L_p) c_k q_{b,k,i} m_{b,k,j}
 * next to logL and (if d_grad is not NULL) the branch gradient, which is then contracted from G
 * (sum_k r_k <Q, G_{b,k}>: this walk leaves the q.(Q m) numerators out).  Costs 1.7-2.0x a
 * b2p_loglik_grad of a default instance.  Such an instance still serves b2p_loglik /
 * b2p_loglik_grad, about a third slower than a default one (one column per thread).  Replaces
 * the 16 K evaluations of b2p_set_gradient_operator; same reference rows: torchtree
 * differentiates these parameters by autograd of the whole recursion
 * (/root/reference/modules/micro.nf:52, /root/reference/plot.Rmd:305).  Sums are added in a
 * fixed order: results are bit-identical from call to call. */
#define B2P_FLAG_OUTER_PRODUCTS 1
int b2p_branch_outer_products(b2p_instance* h, const double* d_bl, double* d_logl, double* d_grad,
                              double* d_outer, void* cuda_stream);

/* Gradients w.r.t. the model from the outer products, on the device (same instance, same
 * branch lengths and model as the b2p_branch_outer_products call that filled d_outer):
 *     d_out[0..15]     dlogL/dQ_ij, row-major, Q's 16 entries taken as free (P = exp(Q r t)):
 *                      V^-T [ sum_{b,k} (V^T G V^-T) o psi(r_k t_b) ] V^T with the divided
 *                      differences psi_ij = (e^{(l_i-l_j) tau} - 1)/(l_i-l_j) of the eigenvalues
 *     d_out[16..19]    dlogL/dpi_i through the root prior only: sum_k G_{root_child,k}[i][i]/pi_i
 *     d_out[20..20+K)  dlogL/dr_k = sum_b t_b <Q, G_{b,k}>
 * The caller chains dQ/dtheta, dpi/dtheta, dr/dtheta of its parameterisation (GTR
 * exchangeabilities, kappa, frequencies, Gamma / Weibull shape) — a few dozen flops.  Two small
 * launches; replaces the 4x4 matrix exponentials b200phylo/paramgrad.py did with torch.
 * root_child: branch id of either child of the root. */
int b2p_model_gradients(b2p_instance* h, const double* d_bl, const double* d_outer, int32_t root_child,
                        double* d_out, void* cuda_stream);

/* Multi-GPU, one process per GPU: pattern shards per rank, results summed
 * over ranks.  With peers attached the final reduction kernel of every evaluation IS the
 * all-reduce: it stores its sums into every peer's receive buffer over NVLink (CUDA IPC
 * peer-mapped memory), exchanges flags and adds the slots in rank order, so d_logl / d_grad
 * come back already summed over all ranks, bit-identical on every rank, with no library
 * collective and no extra launch.  All ranks must then issue their evaluations in the same
 * order (as with any collective).  Set-up, once per instance, on every rank:
 *   b2p_peer_export(h, world, handle64)        allocate the receive buffer, get its IPC handle
 *   <all-gather the 64-byte handles by any means, e.g. torch.distributed>
 *   b2p_peer_attach(h, rank, world, handles)   open the peers' buffers ([world][64] bytes)
 * b2p_peer_status reports whether peers are attached and a nonzero error word if a peer's
 * flag ever failed to arrive within ~4 s (the kernel then gives up instead of hanging).
 * world <= 8 (one NVSwitch domain); 4-state instances only. */
int b2p_peer_export(b2p_instance* h, int32_t world, void* h_handle_out);
int b2p_peer_attach(b2p_instance* h, int32_t rank, int32_t world, const void* h_handles);
int b2p_peer_status(b2p_instance* h, int32_t* attached, int32_t* error_word);

/* Kernel timing for roofline reports: when enabled, the traversal kernel of every
 * evaluation is bracketed by CUDA events on the evaluation's stream;
 * b2p_last_kernel_ms synchronises on the last pair and returns its duration. */
int b2p_set_profiling(b2p_instance* h, int enabled);
int b2p_last_kernel_ms(b2p_instance* h, float* ms_out);
/* Sum of the traversal-kernel durations of the evaluations issued since profiling was
 * enabled (the most recent 1024 at most) and their count; synchronises on them. */
int b2p_kernel_ms_stats(b2p_instance* h, double* total_ms, int32_t* n_launches);

/* Debug / test access: copy what internal non-root node `node` sent to its parent in
 * the last evaluation, m = P_node x_node (x_node = post-order partial, possibly rescaled
 * by a power of two per pattern), as double [P][K][S] into h_out.  BEAGLE: beagleGetPartials. */
int b2p_get_branch_partials(b2p_instance* h, int32_t node, double* h_out);

/* ------------------------------------------------------------------------------------
 * Time trees: the functions the micro benchmark times next to
 * the tree likelihood — ratio_transform, ratio_transform_jacobian, coalescent
 * (/root/reference/plot.Rmd:49-60) — and the pieces the macro model chains in front of it:
 * ReparameterizedTimeTreeModel ratios / root height -> node heights -> branch lengths x
 * strict-clock rate (/root/reference/flu_H3N2/phylotorch-JC69.template:14-58), constant
 * coalescent prior (:75-100).  In torchtree these are GeneralNodeHeightTransform and
 * ConstantCoalescent (Python loops over nodes / a sort); physher has them in C behind
 * "reparam" / "include_jacobian" (/root/reference/flu_H3N2/physher-JC69.template:9,43-44).
 *
 * A b2p_timetree holds one dated topology on one device.  Vectors:
 *   params  [N-1]  ratios of internal nodes N..2N-3 (each in [0,1]), then the root height
 *   heights [N-1]  heights of internal nodes N..2N-2 (time before the most recent tip)
 *   bl      [2N-2] branch lengths by child node id, as b2p_loglik takes them
 * Every evaluation call is one kernel launch on `cuda_stream`, asynchronous, device
 * pointers only (all fp64), no hidden synchronisation; calls on one handle must be issued
 * to one stream at a time (they share scratch).  *_vjp calls return J^T g for an upstream
 * gradient g (what autograd's backward needs). */
typedef struct b2p_timetree b2p_timetree;

/* h_triples as in b2p_set_topology; h_tip_heights [N] >= 0 (sampling dates as heights). */
int b2p_timetree_create(int32_t n_tips, const int32_t* h_triples, const double* h_tip_heights,
                        int32_t device, b2p_timetree** out);
int b2p_timetree_destroy(b2p_timetree* t);
const char* b2p_timetree_last_error(const b2p_timetree* t);
/* tree depth in levels, runs of levels the height kernels walk, 1 if they run out of shared memory */
int b2p_timetree_get_info(const b2p_timetree* t, int32_t* n_levels, int32_t* n_segments, int32_t* in_smem);

/* ratio_transform: h_root = params[N-2]; h_i = b_i + r_i (h_parent(i) - b_i), b_i = latest
 * sampling height below i. */
int b2p_timetree_heights(b2p_timetree* t, const double* d_params, double* d_heights, void* cuda_stream);
/* d_grad_params [N-1] = (d heights / d params)^T d_grad_heights; d_heights = output of the call above. */
int b2p_timetree_heights_vjp(b2p_timetree* t, const double* d_params, const double* d_heights,
                             const double* d_grad_heights, double* d_grad_params, void* cuda_stream);
/* ratio_transform_jacobian: *d_logj = sum_{i != root} log(h_parent(i) - b_i); d_grad_heights [N-1]
 * its gradient w.r.t. the internal heights.  Either output may be NULL. */
int b2p_timetree_log_jacobian(b2p_timetree* t, const double* d_heights, double* d_logj,
                              double* d_grad_heights, void* cuda_stream);
/* bl[c] = rate (h_parent(c) - h_c); d_rate points at one double on the device, NULL = 1. */
int b2p_timetree_branch_lengths(b2p_timetree* t, const double* d_heights, const double* d_rate,
                                double* d_bl, void* cuda_stream);
/* d_grad_heights [N-1] and *d_grad_rate from d_grad_bl [2N-2]; either output may be NULL. */
int b2p_timetree_branch_lengths_vjp(b2p_timetree* t, const double* d_heights, const double* d_rate,
                                    const double* d_grad_bl, double* d_grad_heights,
                                    double* d_grad_rate, void* cuda_stream);
/* coalescent (constant population size theta = *d_theta): log density of the node heights,
 * gradient w.r.t. the internal heights [N-1] and w.r.t. theta.  Outputs may be NULL;
 * d_heights must be 16-byte aligned (any cudaMalloc / torch allocation is). */
int b2p_coalescent_constant(b2p_timetree* t, const double* d_heights, const double* d_theta,
                            double* d_logp, double* d_grad_heights, double* d_grad_theta,
                            void* cuda_stream);

/* The macro benchmark's model density and its gradient in one call
 * (/root/reference/flu_H3N2/phylotorch-JC69.template:11-100: tree likelihood under a strict
 * clock + constant coalescent on the node heights + log-Jacobian of the ratio transform; the
 * ADVI loop evaluates exactly this, value and gradient, once per iteration,
 * /root/reference/modules/macro_flu.nf:133).  Chains ratios/root height -> heights -> branch
 * lengths -> b2p_loglik_grad -> back to d/d(ratios, root height), d/d rate, d/d theta on the
 * device: nine launches on `cuda_stream`, no synchronisation, no host round trip.
 *   like             likelihood instance for the same topology (same device)
 *   d_rate           strict-clock rate (one double), NULL = 1
 *   d_theta          population size; NULL leaves the coalescent term out
 *   include_jacobian add log|det J| of the ratio transform
 *   d_out [4]        total, log-likelihood, coalescent log-density, log-Jacobian (absent terms are
 *                    left untouched and not added)
 *   d_grad_params [N-1], d_grad_rate [1] (may be NULL), d_grad_theta [1] (required with d_theta) */
int b2p_joint_density_grad(b2p_instance* like, b2p_timetree* t, const double* d_params,
                           const double* d_rate, const double* d_theta, int32_t include_jacobian,
                           double* d_out, double* d_grad_params, double* d_grad_rate,
                           double* d_grad_theta, void* cuda_stream);

/* ------------------------------------------------------------------ no-allocation entry point
 * NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid
 * it).  The constant-population coalescent density of b2p_coalescent_constant without a set-up
 * object: the caller passes inputs and outputs only, and the kernel's scratch (per-CTA partial
 * sums + a completion ticket, 8.25 KB) is a file-scope `static __device__` array that every GPU
 * which loads the library holds until the process exits (csrc/b2p_noalloc.cu).  The call makes no
 * cudaMalloc / cudaFree and calls no allocating library routine; it enqueues one 4-byte memset and
 * one kernel on `cuda_stream` and does not synchronise.
 *   device            CUDA device the pointers live on
 *   n_tips            2 .. B2P_NOALLOC_MAX_TIPS (larger: B2P_ERR_UNSUPPORTED with a message)
 *   d_tip_heights [N] tip heights, d_heights [N-1] internal node heights, d_theta [1]
 *   d_logp [1], d_grad_heights [N-1], d_grad_theta [1]: outputs, each may be NULL (not all)
 * Every call on a device shares the one scratch array: calls that OVERLAP on the GPU (different
 * streams without an ordering between them) race — serialise them. */
#define B2P_NOALLOC_MAX_TIPS 16384
int b2p_coalescent_constant_noalloc(int32_t device, int32_t n_tips, const double* d_tip_heights,
                                    const double* d_heights, const double* d_theta, double* d_logp,
                                    double* d_grad_heights, double* d_grad_theta, void* cuda_stream);
/* B2P_NOALLOC_MAX_TIPS as the loaded library was built with. */
int32_t b2p_noalloc_max_tips(void);
/* Message of the calling thread's last failed b2p_coalescent_constant_noalloc call. */
const char* b2p_noalloc_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* B200PHYLO_H */
