/*
 * pyccb200.h - C ABI of libpyccb200.so: the B200 (sm_100a) engine under PyCC's RHF-CCSD
 * amplitude iteration and (T) correction.
 *
 * The reference (jattakumi/pycc) is pure Python and has no FFI; it has two seams, and every
 * entry point below replaces arithmetic that sits behind one of them:
 *
 *   operator seam  pycc/utils.py:141-187   cc_contract.__call__(subscripts, *operands)
 *                  -> pccb_contract()      one fused, strided, batched FP64/FP32 tensor contraction
 *   solver seam    pycc/ccwfn.py:234-329   solve_cc: amplitude update, norm, energy
 *                  pycc/ccwfn.py:408-410   build_tau;  :621 pair symmetrisation of r2
 *                  pycc/utils.py:6-139     helper_diis (error vectors, B matrix, extrapolation)
 *                  pycc/cctriples.py:68-83,106-152   t3d_ijk + the t_tjl energy reduction
 *                  -> the pccb_* element-wise / reduction entry points
 *
 * Conventions
 *   - plain C: pointers, sizes, scalars; no C++ or torch types cross the boundary.
 *   - every pointer named *_dev is DEVICE memory owned by the caller (PyTorch tensors on the
 *     Python side); the library allocates nothing persistent except small per-context scratch.
 *   - all work is enqueued on the context's CUDA stream (pccb_set_stream); entry points that
 *     return a scalar to the host say so and synchronise that stream.
 *   - return value 0 = OK, negative = error; pccb_last_error() gives the message.
 *   - all reductions are fixed-order (bitwise reproducible run to run).
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef PYCCB200_H
#define PYCCB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PCCB_ABI_VERSION 1
#define PCCB_MAX_SEGMENTS 4

enum { PCCB_OK = 0, PCCB_ERR_ARG = -1, PCCB_ERR_CUDA = -2, PCCB_ERR_NODEV = -3, PCCB_ERR_WORKSPACE = -4 };
enum { PCCB_F64 = 0, PCCB_F32 = 1 };
/* contraction kernel selector: the tensor-core tile kernel is the product path; the naive
 * one-thread-per-element kernel exists so GPU tests can cross-check it on the same descriptor */
enum { PCCB_IMPL_TENSOR = 0, PCCB_IMPL_NAIVE = 1 };

typedef struct pccb_ctx pccb_ctx;

/* Two-level compound index: element offset of logical index i is
 *     (i / div) * s_hi + (i % div) * s_lo          (offsets and strides in ELEMENTS).
 * A flat index has div >= extent (s_hi unused).  This is what lets one GEMM tile walk an
 * einsum index group such as (i,a) of t2[i,m,a,e] without a transpose ever landing in HBM. */
typedef struct pccb_cidx {
    int64_t div;
    int64_t s_hi;
    int64_t s_lo;
} pccb_cidx;

/* One K-segment:  sum_k A[am(m) + ak(k) + batch*a_batch] * B[bn(n) + bk(k) + batch*b_batch] * scale.
 * Several segments accumulate into the same output tile (K-concatenation of structurally equal
 * einsum terms), so e.g. the 12 terms of the (T) W intermediate need 3 launches, not 12. */
typedef struct pccb_segment {
    const void* A_dev;
    const void* B_dev;
    pccb_cidx am, ak, bn, bk;
    int64_t K;
    int64_t a_batch, b_batch;
    double scale;
} pccb_segment;

/* C[cm(m)+cn(n)+b*c_batch] = alpha * sum_segments(...) + beta * C[...] + gamma * D[dm(m)+dn(n)+b*d_batch]
 * (D_dev may be NULL; then gamma is ignored).  Replaces one or more cc_contract calls
 * (pycc/utils.py:162-187) including the permuted accumulation into the target array. */
typedef struct pccb_contraction {
    int64_t M, N, batch;
    int32_t nseg;
    int32_t dtype; /* PCCB_F64 (DMMA) or PCCB_F32 (3xTF32 tensor-core products, FP32 accumulation) */
    pccb_segment seg[PCCB_MAX_SEGMENTS];
    void* C_dev;
    pccb_cidx cm, cn;
    int64_t c_batch;
    const void* D_dev;
    pccb_cidx dm, dn;
    int64_t d_batch;
    double alpha, beta, gamma;
} pccb_contraction;

/* ---- context ------------------------------------------------------------------------- */
int pccb_abi_version(void);
int pccb_device_count(void); /* 0 when no CUDA device / driver */
int pccb_create(int device, pccb_ctx** out);
int pccb_destroy(pccb_ctx* ctx);
const char* pccb_last_error(pccb_ctx* ctx); /* ctx may be NULL: last error of a failed create */
int pccb_set_stream(pccb_ctx* ctx, void* cuda_stream);
/* split-K partial tiles live here; caller-owned device buffer (>= 64 MiB recommended) */
int pccb_set_workspace(pccb_ctx* ctx, void* ws_dev, size_t bytes);
int pccb_set_contract_impl(pccb_ctx* ctx, int impl);
int pccb_launch_count(pccb_ctx* ctx, int64_t* n_out); /* kernels launched so far by this ctx */
int pccb_synchronize(pccb_ctx* ctx);

/* ---- operator seam: cc_contract (pycc/utils.py:162-187) --------------------------------- */
int pccb_contract(pccb_ctx* ctx, const pccb_contraction* desc);

/* Mixed-precision mode (precision='SP', pycc/ccwfn.py:189-196), tcgen05 path: an FP32 operand tensor
 * x can be given a residual plane lo = x - trunc_tf32(x) (pccb_split_lo writes it; same layout as x).
 * Once both operands of a plain-matrix FP32 contraction have registered planes, pccb_contract runs it
 * as 3xTF32 on the 5th-generation tensor cores (A B ~= A B + A_lo B + A B_lo, FP32 accumulation in
 * tensor memory); otherwise the mma.sync 3xTF32 kernel is used.  Registration is by address range:
 * views into a registered tensor are covered. */
int pccb_split_lo(pccb_ctx* ctx, const void* x_dev, void* lo_dev, int64_t n);
int pccb_register_lo_plane(pccb_ctx* ctx, const void* x_dev, const void* lo_dev, int64_t n);
int pccb_clear_lo_planes(pccb_ctx* ctx);
/* forget the plane registered for the tensor whose base address is x_dev (no-op when there is none) */
int pccb_unregister_lo_plane(pccb_ctx* ctx, const void* x_dev);

/* ---- solver seam: element-wise and reduction pieces ------------------------------------- */
/* out[i,j,a,b] = c0*X[i,j,a,b] + c1*X[i,j,b,a] + c2*u[i,a]*w[j,b]   (u_dev/w_dev may be NULL)
 * build_tau (pycc/ccwfn.py:408-410), 2*t2 - t2.swapaxes(2,3) (:554,556), t2 - t2.swapaxes(2,3)
 * (:609) and L = 2*ERI - ERI.swapaxes(2,3) on the oovv block (pycc/hamiltonian.py:39). */
int pccb_pair_combine(pccb_ctx* ctx, int dtype, void* out_dev, const void* X_dev, double c0, double c1,
                      const void* u_dev, const void* w_dev, double c2, int64_t no, int64_t nv);

/* r2[i,j,a,b] += r2[j,i,b,a] in place (pycc/ccwfn.py:621). */
int pccb_pair_symmetrize(pccb_ctx* ctx, int dtype, void* r2_dev, int64_t no, int64_t nv);

/* out[d0,d1,d2,d3] (C-contiguous) = alpha * in[strided view] + beta * out.  Used once at
 * set-up to lay the ERI blocks out in the orders the contractions consume. */
int pccb_permute4(pccb_ctx* ctx, int dtype, void* out_dev, const void* in_dev, const int64_t shape[4],
                  const int64_t in_strides[4], double alpha, double beta);

/* t1 += r1/Dia; t2 += r2/Dijab with D from the Fock diagonal eps (occupied first);
 * *rms2_host = sum (r1/Dia)^2 + sum (r2/Dijab)^2  (pycc/ccwfn.py:287-290).
 * If symmetrize != 0, r2 is taken as r2[i,j,a,b] + r2[j,i,b,a] on the fly (ccwfn.py:621) and
 * is left untouched.  Host-returning: synchronises the stream. */
int pccb_update_amplitudes(pccb_ctx* ctx, int dtype, void* t1_dev, void* t2_dev, const void* r1_dev,
                           const void* r2_dev, const void* eps_dev, int64_t no, int64_t nv, int symmetrize,
                           double* rms2_host);

/* Fused form of the two calls around this comment, for the (i,j) blocks ij = i*no+j in [ij_lo, ij_hi)
 * of t2 (the whole range on one GPU; one rank's share when the update is split over ranks):
 * t1 += r1/Dia (always, all of it), t2[ij] += (r2[ijab] (+ r2[jiba])) / Dijab, and the energy of the NEW
 * amplitudes over those blocks.  out_host[0] = sum of squared increments, out_host[1] = energy part;
 * the t1 increment norm and the singles energy 2 sum F_ia t_ia are included only when count_t1 != 0, so
 * that the ranks' results add up to pycc/ccwfn.py:287-296.  Host-returning. */
int pccb_update_energy(pccb_ctx* ctx, int dtype, void* t1_dev, void* t2_dev, const void* r1_dev, const void* r2_dev,
                       const void* eps_dev, const void* fov_dev, int64_t ld_fov, const void* oovv_dev, int64_t no,
                       int64_t nv, int symmetrize, int64_t ij_lo, int64_t ij_hi, int count_t1, double* out_host);

/* One pass over t2 that writes every amplitude-dependent o^2 v^2 operand of the planned contractions
 * (any pointer may be NULL): tau = t2 + t1 t1, tauh = t2 + t1 t1 / 2, tauW = t2 / 2 + t1 t1
 * (build_tau, pycc/ccwfn.py:408-410), t2s = 2 t2 - t2.swapaxes(2,3) (:554), t11 = t1 t1 (CC2), and the
 * plain-matrix layouts t2sP[i,a,m,e] = t2s[i,m,a,e], t2P[i,a,m,e] = t2[i,m,a,e], t2Q[j,a,m,e] = t2[m,j,a,e],
 * tauWPn[j,B,n,f] = -tauW[j,n,f,B], t2Rh[j,B,n,f] = t2[n,j,f,B] / 2 with B restricted to [vlo, vlo+vl). */
typedef struct pccb_prepare_outputs {
    void *tau, *tauh, *tauW, *t2s, *t11, *t2sP, *t2P, *t2Q, *tauWPn, *t2Rh;
} pccb_prepare_outputs;
int pccb_prepare_amplitudes(pccb_ctx* ctx, int dtype, const void* t1_dev, const void* t2_dev, int64_t no, int64_t nv,
                            int64_t vlo, int64_t vl, const pccb_prepare_outputs* out);

/* E = 2 sum F_ia t_ia + sum (t2_ijab + t_ia t_jb) (2<ij|ab> - <ij|ba>)  (pycc/ccwfn.py:717-718).
 * fov_dev is the (no,nv) occupied-virtual Fock block, row stride ld_fov.  Host-returning. */
int pccb_cc_energy(pccb_ctx* ctx, int dtype, const void* fov_dev, int64_t ld_fov, const void* t1_dev,
                   const void* t2_dev, const void* oovv_dev, int64_t no, int64_t nv, double* e_host);

/* pccb_cc_energy without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving
 * loops, harnesses that forbid it).  Same arithmetic and summation order as pccb_cc_energy, but
 *   - the per-CTA partial sums live in a file-scope static __device__ array of
 *     pccb_cc_energy_static_max_partials() = 131072 doubles (1 MiB; every GPU that loads the library holds
 *     it until the process exits); problems with ceil(nv/32)^2 * no^2 above that are rejected with
 *     PCCB_ERR_ARG and a message (e.g. no=30 nv=250 needs 57600, no=40 nv=400 needs 270400: too large);
 *   - the energy is written to e_dev, a DEVICE double the caller owns; the call is asynchronous on the
 *     ctx stream and calls no cudaMalloc / cudaFree / cudaMallocHost and no allocating library routine.
 * Every call shares the one array: calls that overlap on the GPU (two streams, concurrent graph branches,
 * two contexts on one device) race on it; serialise them.  Nothing needs clearing between calls (each call
 * overwrites every slot it sums). */
int64_t pccb_cc_energy_static_max_partials(void);
int pccb_cc_energy_static(pccb_ctx* ctx, int dtype, const void* fov_dev, int64_t ld_fov, const void* t1_dev,
                          const void* t2_dev, const void* oovv_dev, int64_t no, int64_t nv, double* e_dev);

/* DIIS building blocks (pycc/utils.py:26-46, 108-133); accumulation is FP64 for both dtypes. */
int pccb_dot(pccb_ctx* ctx, int dtype, const void* x_dev, const void* y_dev, int64_t n, double* out_host);
/* out_host[k] = <x, ys[k]>, k < m <= 16, in ONE pass over x and one host synchronisation: the new row of the DIIS
 * Gram matrix (pycc/utils.py:108-112 fills B[a,b] with one np.dot per entry). */
int pccb_dot_many(pccb_ctx* ctx, int dtype, const void* x_dev, const void* const* ys_dev_hostarray, int m, int64_t n,
                  double* out_host);
int pccb_axpby(pccb_ctx* ctx, int dtype, void* out_dev, double a, const void* x_dev, double b, const void* y_dev,
               int64_t n); /* out = a*x + b*y (y_dev may be NULL) */
int pccb_lincomb(pccb_ctx* ctx, int dtype, void* out_dev, const void* const* vecs_dev_hostarray,
                 const double* coef_host, int nvec, int64_t n); /* out = sum_k coef[k]*vec[k] */

/* ---- (T): disconnected triples + Lee-Rendell energy reduction -----------------------------
 * For `ntrip` occupied triplets (i>=j>=k, device array ijk_dev[3*ntrip]) whose connected W[a,b,c]
 * (pycc/cctriples.py:10-22, built by 3 pccb_contract launches per batch) sit in
 * w_dev[t][a][b][c]: forms V = W + t3d_ijk (cctriples.py:69-74), the X/Y/Z combinations and
 * the restricted sum over a>=b>=c with the occupation factor and denominator
 * (cctriples.py:124-150).  et_dev[t] (DEVICE, always double) receives each triplet's energy.
 * Asynchronous: the caller sums et_dev after pccb_synchronize(). */
int pccb_triples_energy(pccb_ctx* ctx, int dtype, const void* w_dev, const int32_t* ijk_dev, int ntrip,
                        const void* t1_dev, const void* t2_dev, const void* oovv_dev, const void* fov_dev,
                        int64_t ld_fov, const void* eps_dev, int64_t no, int64_t nv, double* et_dev);

/* ---- CC3 (next row): t3 = W / D for k = k0 .. k0+nk-1 of the occupied pair (i, j) and the
 * two combinations contracted into r1 / r2 (pycc/ccwfn.py:387-398, pycc/cctriples.py:24-33):
 *   a1[k][a][b][c] = t3[abc] - t3[cba],   a2[k][a][b][c] = 2 t3[abc] - t3[acb] - t3[cba].
 * w_dev holds the connected triples W[k][a][b][c] built by pccb_contract from the T1-dressed
 * integrals (ccwfn.py:625-710).  Asynchronous. */
int pccb_t3_combine(pccb_ctx* ctx, int dtype, const void* w_dev, int nk, int k0, int i, int j, const void* eps_dev,
                    int64_t no, int64_t nv, void* a1_dev, void* a2_dev);

/* ---- (T) contributions to the Lambda equations and densities (next row; pycc/ccwfn.py:721-801
 * t3_density): for the occupied pair (i, j) and k = k0 .. k0+nk-1, from the connected w_dev[k][a][b][c] and the
 * disconnected dd_dev[k][a][b][c] triples (both without denominator; pycc/cctriples.py:8-35, 68-83), the seven
 * combinations the density / Lambda contractions consume, out7 = {M, A1, A2, B1, Z3, P, Q} (any may be NULL):
 *   M = W/D, N = Dd/D, X(T) = 8T - 4T[bac] - 4T[acb] - 4T[cba] + 2T[cab] + 2T[bca],
 *   A1 = M - M[cba], A2 = 2M - M[acb] - M[cba], B1 = 2(M - M[bac]), Z3 = 2(M - M[acb]) - (M[bac] - M[bca]),
 *   P = 2X(M) + X(N), Q = X(M) + X(N).  Only the permutations the non-NULL outputs consume are read (dd_dev is read
 * for P / Q only and may alias w_dev otherwise: the CC3 Lambda engine asks for M or A1 alone, pycc/cclambda.py:133,
 * 193, 206).  Asynchronous. */
int pccb_t3_density_combine(pccb_ctx* ctx, int dtype, const void* w_dev, const void* dd_dev, int nk, int k0, int i, int j,
                            const void* eps_dev, int64_t no, int64_t nv, void* const* out7_dev_hostarray);

#ifdef __cplusplus
}
#endif
#endif /* PYCCB200_H */
