// (T) energy reduction: disconnected triples + Lee-Rendell closed-shell energy expression,
// evaluated per occupied triplet from the connected W[a,b,c] that pccb_contract produced
// (pycc/cctriples.py:68-83 t3d_ijk, :124-150 body of t_tjl).
//
// The expression needs W and V at all six permutations of (a,b,c).  A CTA owns the 8x8 (a,b)
// brick pair (TA >= TB) and walks TC <= TB; for each canonical brick triple it stages the six
// permuted 8x8x8 bricks of W in shared memory with coalesced 64-byte runs, so W is read from
// HBM/L2 exactly once per triplet (8*v^3 bytes: HBM-roofline work, ~1 % of the W build).
// Reductions are fixed-order: warp shuffle -> block -> one partial per CTA -> one block per
// triplet.
#include "common.cuh"
#include <cstdlib>

namespace {

constexpr int TB_ = 8;                 // brick edge
constexpr int BRICK = TB_ * TB_ * TB_; // dense [8][8][8], XOR-swizzled (brick_off)

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

template <typename T>
struct TriplesArgs {
    const T* W;           // [ntrip][nv][nv][nv]
    const int* ijk;       // device [3*ntrip]
    const T* t1;          // [no][nv]
    const T* t2;          // [no][no][nv][nv]
    const T* g;           // oovv
    const T* fov;         // F[o,v] with row stride ld_fov
    const T* eps;         // Fock diagonal, occupied first
    long long ld_fov;
    int no, nv, nb, ntrip;
    double* partials;     // [ntrip][nb*(nb+1)/2]
};

// element (x0+x, y0+y, z0+z) of W for this thread's (x,y,z), zero beyond nv
template <typename T>
__device__ __forceinline__ double fetch_brick(const T* W, int nv, int x0, int y0, int z0, int tid) {
    const int z = tid & 7, y = (tid >> 3) & 7, x = tid >> 6;
    const int gx = x0 + x, gy = y0 + y, gz = z0 + z;
    double v = 0.0;
    if (gx < nv && gy < nv && gz < nv) v = (double)__ldg(&W[((long long)gx * nv + gy) * nv + gz]);
    return v;
}
// Shared-memory position of brick element (x,y,z).  A brick is written in the natural thread order (z fastest) and
// read at one of the six permutations of the thread's (la,lb,lc); with 64-bit accesses a half-warp spans 8 values of
// one coordinate and 2 of another, and has to land in 16 distinct 8-byte banks.  Dense layout with two XORs:
//   bank bit 3 = y0 ^ x0,  bank bits 2..0 = z ^ (x0 | (y2 y1 ^ x2 x1) << 1)
// is a bijection of the 4 varying bits for the store and for all six read patterns (0 conflicts; the padded
// [8][8][9] layout it replaces took 832 instead of 384 wavefronts per brick triple).
__device__ __forceinline__ int brick_off(int x, int y, int z) {
    return (x << 6) + ((y ^ (x & 1)) << 3) + (z ^ ((x & 1) | ((((y >> 1) ^ (x >> 1)) & 3) << 1)));
}
__device__ __forceinline__ void store_brick(double* s, double v, int tid) {
    const int z = tid & 7, y = (tid >> 3) & 7, x = tid >> 6;
    s[brick_off(x, y, z)] = v;
}

__device__ __forceinline__ double brick_at(const double* s, int x, int y, int z) {
    return s[brick_off(x, y, z)];
}

// Shared-memory staging of the disconnected part: for the six v x v matrices
//   m = 0..5 : <ij|..>, <ik|..>, <jk|..>, t2[ij], t2[ik], t2[jk]
// the expression reads element (x,y) at every ordered pair of the three brick ranges, i.e. six 8x8
// blocks per matrix: AB, BA (fixed for the CTA) and AC, CA, BC, CB (per c-tile).  Each block is
// fetched once per CTA / c-tile with 64-byte runs (3 loads per thread per c-tile) instead of the 36
// scattered global loads per element the direct evaluation needs.
constexpr int GBLK = TB_ * TB_;         // dense [8][8], column XOR row (goff): conflict-free for the natural
                                        // store and for (a,b), (a,c), (b,c) indexed reads in either order
__device__ __forceinline__ int goff(int r, int s) { return (r << 3) + (s ^ r); }

template <typename T>
__global__ void __launch_bounds__(512) triples_energy_kernel(const __grid_constant__ TriplesArgs<T> p) {
    __shared__ double sW[6][BRICK];
    __shared__ double sG[36][GBLK];       // [matrix * 6 + pair type][8][8 ^ row]
    __shared__ double sred[16];
    const int TBk = blockIdx.x, TAk = blockIdx.y, t = blockIdx.z;
    if (TBk > TAk) return;
    const int tid = threadIdx.x;
    const int lc = tid & 7, lb = (tid >> 3) & 7, la = tid >> 6;
    const int nv = p.nv, no = p.no;
    const int i = p.ijk[3 * t], j = p.ijk[3 * t + 1], k = p.ijk[3 * t + 2];
    const T* W = p.W + (long long)t * nv * nv * nv;
    const long long nvv = (long long)nv * nv;
    // matrix m = 0..5 : <ij|..>, <ik|..>, <jk|..>, t2[ij], t2[ik], t2[jk];  vector m (its partner in the
    // disconnected part): t1[k], t1[j], t1[i], f[k], f[j], f[i]
    auto matp = [&](int m) -> const T* {
        const int mm = m >= 3 ? m - 3 : m;
        const int x = mm == 2 ? j : i, y = mm == 0 ? j : k;
        return (m >= 3 ? p.t2 : p.g) + ((long long)x * no + y) * nvv;
    };
    auto vecp = [&](int m) -> const T* {
        const int mm = m >= 3 ? m - 3 : m;
        const int x = mm == 0 ? k : (mm == 1 ? j : i);
        return m >= 3 ? p.fov + (long long)x * p.ld_fov : p.t1 + (long long)x * nv;
    };
    const double eocc = (double)p.eps[i] + (double)p.eps[j] + (double)p.eps[k];
    const double occ = 2.0 - (double)((i == j) + (i == k) + (j == k));

    const int A0 = TAk * TB_, B0 = TBk * TB_;
    const int a = A0 + la, b = B0 + lb;
    double esum = 0.0;

    // one 8x8 block element of matrix m at (r0 + r, c0 + s), zero beyond nv
    auto gfetch = [&](int m, int r0, int c0, int e) -> double {
        const int r = r0 + (e >> 3), c = c0 + (e & 7);
        return (r < nv && c < nv) ? (double)__ldg(matp(m) + (long long)r * nv + c) : 0.0;
    };
    // fixed blocks AB / BA: 12 blocks x 64 elements
    for (int e = tid; e < 12 * 64; e += 512) {
        const int q = e >> 6, m = q >> 1, pt = q & 1;
        sG[m * 6 + pt][goff((e & 63) >> 3, e & 7)] = pt == 0 ? gfetch(m, A0, B0, e & 63) : gfetch(m, B0, A0, e & 63);
    }
    // per-thread vector values at a and b (position 0 and 1); the c values are read per c-tile
    double va[6], vb[6];
#pragma unroll
    for (int m = 0; m < 6; ++m) {
        va[m] = a < nv ? (double)vecp(m)[a] : 0.0;
        vb[m] = b < nv ? (double)vecp(m)[b] : 0.0;
    }
    const double eab = (a < nv ? (double)p.eps[no + a] : 0.0) + (b < nv ? (double)p.eps[no + b] : 0.0);

    // the six permuted bricks of W and the 24 c-dependent matrix blocks of the NEXT c-tile are fetched
    // into registers while the current one is being consumed from shared memory
    double nxt[6], gnx[3];
    auto fetch = [&](int C0) {
        nxt[0] = fetch_brick(W, nv, A0, B0, C0, tid);   // [a][b][c]
        nxt[1] = fetch_brick(W, nv, A0, C0, B0, tid);   // [a][c][b]
        nxt[2] = fetch_brick(W, nv, B0, A0, C0, tid);   // [b][a][c]
        nxt[3] = fetch_brick(W, nv, B0, C0, A0, tid);   // [b][c][a]
        nxt[4] = fetch_brick(W, nv, C0, A0, B0, tid);   // [c][a][b]
        nxt[5] = fetch_brick(W, nv, C0, B0, A0, tid);   // [c][b][a]
#pragma unroll
        for (int u = 0; u < 3; ++u) {
            const int e = tid + 512 * u, q = e >> 6, m = q >> 2, pt = q & 3;    // pt: AC, CA, BC, CB
            const int r0 = pt == 0 ? A0 : (pt == 2 ? B0 : C0), c0 = pt == 1 ? A0 : (pt == 3 ? B0 : C0);
            gnx[u] = gfetch(m, r0, c0, e & 63);
        }
    };
    fetch(0);
    for (int TCk = 0; TCk <= TBk; ++TCk) {
        const int C0 = TCk * TB_;
        const int c = C0 + lc;
        __syncthreads();
#pragma unroll
        for (int q = 0; q < 6; ++q) store_brick(sW[q], nxt[q], tid);
#pragma unroll
        for (int u = 0; u < 3; ++u) {
            const int e = tid + 512 * u, q = e >> 6, m = q >> 2, pt = q & 3;
            sG[m * 6 + 2 + pt][goff((e & 63) >> 3, e & 7)] = gnx[u];
        }
        __syncthreads();
        if (TCk < TBk) fetch(C0 + TB_);
        if (a < nv && b < nv && c < nv && a >= b && b >= c) {
            const double Wabc = brick_at(sW[0], la, lb, lc);
            const double Wacb = brick_at(sW[1], la, lc, lb);
            const double Wbac = brick_at(sW[2], lb, la, lc);
            const double Wbca = brick_at(sW[3], lb, lc, la);
            const double Wcab = brick_at(sW[4], lc, la, lb);
            const double Wcba = brick_at(sW[5], lc, lb, la);
            double vc[6];
#pragma unroll
            for (int m = 0; m < 6; ++m) vc[m] = (double)__ldg(vecp(m) + c);
            // disconnected part for the virtual indices at positions (px,py,pz) of (a,b,c) (cctriples.py:69-74):
            //   g_ij[x,y] t1k[z] + g_ik[x,z] t1j[y] + g_jk[y,z] t1i[x] + t_ij[x,y] fk[z] + t_ik[x,z] fj[y] + t_jk[y,z] fi[x]
            auto blk = [&](int m, int px, int py) -> double {
                const int pt = px == 0 ? (py == 1 ? 0 : 2) : (px == 1 ? (py == 0 ? 1 : 4) : (py == 0 ? 3 : 5));
                return sG[m * 6 + pt][goff(px == 0 ? la : (px == 1 ? lb : lc), py == 0 ? la : (py == 1 ? lb : lc))];
            };
            auto vat = [&](int m, int pos) -> double { return pos == 0 ? va[m] : (pos == 1 ? vb[m] : vc[m]); };
            auto dis = [&](int px, int py, int pz) -> double {
                return blk(0, px, py) * vat(0, pz) + blk(1, px, pz) * vat(1, py) + blk(2, py, pz) * vat(2, px) +
                       blk(3, px, py) * vat(3, pz) + blk(4, px, pz) * vat(4, py) + blk(5, py, pz) * vat(5, px);
            };
            const int ndel = (a == b) + (a == c) + (b == c);                  // 0, 1 or 3
            const double sinv = ndel == 0 ? 1.0 : (ndel == 1 ? 0.5 : 0.25);    // 1 / (1 + ndel)
            const double Vabc = (Wabc + dis(0, 1, 2)) * sinv;
            const double Vacb = (Wacb + dis(0, 2, 1)) * sinv;
            const double Vbac = (Wbac + dis(1, 0, 2)) * sinv;
            const double Vbca = (Wbca + dis(1, 2, 0)) * sinv;
            const double Vcab = (Wcab + dis(2, 0, 1)) * sinv;
            const double Vcba = (Wcba + dis(2, 1, 0)) * sinv;
            const double X = Wabc * Vabc + Wacb * Vacb + Wbac * Vbac + Wbca * Vbca + Wcab * Vcab + Wcba * Vcba;
            const double Y = Vabc + Vbca + Vcab;
            const double Z = Vacb + Vbac + Vcba;
            const double Wc = Wabc + Wbca + Wcab;
            const double Wa = Wacb + Wbac + Wcba;
            const double denom = eocc - eab - (double)p.eps[no + c];
            esum += ((Y - 2.0 * Z) * Wc + (Z - 2.0 * Y) * Wa + 3.0 * X) * occ / denom;
        }
    }
    esum = warp_sum(esum);
    if ((tid & 31) == 0) sred[tid >> 5] = esum;
    __syncthreads();
    if (tid == 0) {
        double r = 0.0;
#pragma unroll
        for (int w = 0; w < 16; ++w) r += sred[w];
        const int npair = p.nb * (p.nb + 1) / 2;
        p.partials[(long long)t * npair + TAk * (TAk + 1) / 2 + TBk] = r;
    }
}

__global__ void sum_rows_kernel(const double* __restrict__ partials, int n, double* __restrict__ out) {
    __shared__ double sh[8];
    const double* row = partials + (long long)blockIdx.x * n;
    double v = 0.0;
    for (int q = threadIdx.x; q < n; q += blockDim.x) v += row[q];
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        double r = 0.0;
        for (int w = 0; w < 8; ++w) r += sh[w];
        out[blockIdx.x] = r;
    }
}


// ---- CC3: amplitudes t3 = W / D for a batch of k and the two spin-adapted combinations the
// residual contractions need (pycc/ccwfn.py:387-398):
//   A1[k][a][b][c] = t3[abc] - t3[cba]                 (t3 - t3.swapaxes(0,2))
//   A2[k][a][b][c] = 2 t3[abc] - t3[acb] - t3[cba]     (2 t3 - t3.swapaxes(1,2) - t3.swapaxes(0,2))
// D^{ijk}_{abc} (pycc/cctriples.py:24-33) is invariant under permutations of (a,b,c).
template <typename T>
__global__ void t3_combine_kernel(const T* __restrict__ W, int nk, int k0, int i, int j, const T* __restrict__ eps,
                                  int no, int nv, T* __restrict__ A1, T* __restrict__ A2) {
    const long long v2 = (long long)nv * nv, v3 = v2 * nv, total = v3 * nk;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int kk = (int)(idx / v3);
        const long long r = idx - (long long)kk * v3;
        const int a = (int)(r / v2), b = (int)((r / nv) % nv), c = (int)(r % nv);
        const T* Wk = W + (long long)kk * v3;
        const double d = (double)eps[i] + (double)eps[j] + (double)eps[k0 + kk] - (double)eps[no + a] -
                         (double)eps[no + b] - (double)eps[no + c];
        const double abc = (double)Wk[r], cba = (double)Wk[c * v2 + (long long)b * nv + a],
                     acb = (double)Wk[a * v2 + (long long)c * nv + b];
        A1[idx] = (T)((abc - cba) / d);
        A2[idx] = (T)((2.0 * abc - acb - cba) / d);
    }
}

// ---- (T) densities / Lambda source terms (pycc/ccwfn.py:721-801, t3_density): for the occupied pair (i, j) and
// k = k0 .. k0+nk-1, from the connected W[k][abc] and disconnected Dd[k][abc] (both WITHOUT denominator):
//   M = W / D,  N = Dd / D                                            (t3c_ijk / t3d_ijk with WithDenom=True)
//   X(T)[abc] = 8 T[abc] - 4 T[bac] - 4 T[acb] - 4 T[cba] + 2 T[cab] + 2 T[bca]            (ccwfn.py:754-755)
//   A1 = M - M[cba]                 A2 = 2 M - M[acb] - M[cba]         (:758-760, 766)
//   B1 = 2 (M - M[bac])             Z3 = 2 (M - M[acb]) - (M[bac] - M[bca])               (:775, 769)
//   P  = 2 X(M) + X(N)              Q  = X(M) + X(N)                   (:771-772, 777-778; :763, 790)
// D^{ijk}_{abc} is invariant under permutations of (a, b, c).  Any output pointer may be null.
template <typename T>
struct T3DensityOut { T *M, *A1, *A2, *B1, *Z3, *P, *Q; };

template <typename T>
__global__ void t3_density_combine_kernel(const T* __restrict__ W, const T* __restrict__ Dd, int nk, int k0, int i, int j,
                                          const T* __restrict__ eps, int no, int nv,
                                          const __grid_constant__ T3DensityOut<T> o) {
    const long long v2 = (long long)nv * nv, v3 = v2 * nv, total = v3 * nk;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const int kk = (int)(idx / v3);
        const long long r = idx - (long long)kk * v3;
        const int a = (int)(r / v2), b = (int)((r / nv) % nv), c = (int)(r % nv);
        const T* Wk = W + (long long)kk * v3;
        const T* Dk = Dd + (long long)kk * v3;
        const double d = (double)eps[i] + (double)eps[j] + (double)eps[k0 + kk] - (double)eps[no + a] -
                         (double)eps[no + b] - (double)eps[no + c];
        const double rd = 1.0 / d;      // one division per element; the twelve permuted values are scaled by it
        auto at = [&](const T* X, int x, int y, int z) { return (double)X[x * v2 + (long long)y * nv + z] * rd; };
        // only the permutations the requested outputs consume are read (the transposed ones are strided accesses:
        // a caller that wants M alone - the CC3 Lambda engine - pays one coalesced read); with P / Q requested all
        // twelve loads are issued unconditionally, back to back
        double mabc, mbac, macb, mcba, mcab, mbca, xn = 0.0;
        if (o.P || o.Q) {
            mabc = at(Wk, a, b, c); mbac = at(Wk, b, a, c); macb = at(Wk, a, c, b); mcba = at(Wk, c, b, a);
            mcab = at(Wk, c, a, b); mbca = at(Wk, b, c, a);
            xn = 8.0 * at(Dk, a, b, c) - 4.0 * at(Dk, b, a, c) - 4.0 * at(Dk, a, c, b) - 4.0 * at(Dk, c, b, a) +
                 2.0 * at(Dk, c, a, b) + 2.0 * at(Dk, b, c, a);
        } else {
            mabc = at(Wk, a, b, c);
            mbac = (o.B1 || o.Z3) ? at(Wk, b, a, c) : 0.0;
            macb = (o.A2 || o.Z3) ? at(Wk, a, c, b) : 0.0;
            mcba = (o.A1 || o.A2) ? at(Wk, c, b, a) : 0.0;
            mbca = o.Z3 ? at(Wk, b, c, a) : 0.0;
            mcab = 0.0;
        }
        const double xm = 8.0 * mabc - 4.0 * mbac - 4.0 * macb - 4.0 * mcba + 2.0 * mcab + 2.0 * mbca;
        if (o.M) o.M[idx] = (T)mabc;
        if (o.A1) o.A1[idx] = (T)(mabc - mcba);
        if (o.A2) o.A2[idx] = (T)(2.0 * mabc - macb - mcba);
        if (o.B1) o.B1[idx] = (T)(2.0 * (mabc - mbac));
        if (o.Z3) o.Z3[idx] = (T)(2.0 * (mabc - macb) - (mbac - mbca));
        if (o.P) o.P[idx] = (T)(2.0 * xm + xn);
        if (o.Q) o.Q[idx] = (T)(xm + xn);
    }
}

// Brick form of the two combine kernels (the default): one CTA per 8x8x8 brick (TA, TB, TC) of one k.  The permuted
// values M[bac], M[acb], ... of the CTA's elements live in the bricks (TB,TA,TC), (TA,TC,TB), ...: each is fetched with
// coalesced 64-byte runs and transposed through the XOR-swizzled shared-memory layout of the energy kernel, instead
// of the element-wise kernels' strided 8-byte reads (one useful double per 32-byte sector).  Outputs are written at the
// element's own position, coalesced.  Only the bricks the requested outputs consume are fetched.
template <typename T>
__global__ void __launch_bounds__(512) t3_combine_brick_kernel(const T* __restrict__ W, const T* __restrict__ Dd, int k0,
                                                               int i, int j, const T* __restrict__ eps, int no, int nv,
                                                               int nb, const __grid_constant__ T3DensityOut<T> o) {
    __shared__ double sW[5][BRICK];
    const int tid = threadIdx.x;
    const int lc = tid & 7, lb = (tid >> 3) & 7, la = tid >> 6;
    const int bx = blockIdx.x, kk = blockIdx.y;
    const int TC = bx % nb, TB = (bx / nb) % nb, TA = bx / (nb * nb);
    const int A0 = TA * TB_, B0 = TB * TB_, C0 = TC * TB_;
    const int a = A0 + la, b = B0 + lb, c = C0 + lc;
    const long long v2 = (long long)nv * nv, v3 = v2 * nv;
    const T* Wk = W + (long long)kk * v3;
    const bool need_x = o.P || o.Q;
    const bool n_bac = need_x || o.B1 || o.Z3, n_acb = need_x || o.A2 || o.Z3, n_cba = need_x || o.A1 || o.A2,
               n_bca = need_x || o.Z3;
    // this thread's element of the permuted bricks (brick-local position (la, lb, lc) of each)
    auto gather = [&](const T* X, double& xabc, double& xacb, double& xbac, double& xbca, double& xcab, double& xcba) {
        xabc = fetch_brick(X, nv, A0, B0, C0, tid);
        const double r1 = n_acb ? fetch_brick(X, nv, A0, C0, B0, tid) : 0.0;     // [a][c][b]
        const double r2 = n_bac ? fetch_brick(X, nv, B0, A0, C0, tid) : 0.0;     // [b][a][c]
        const double r3 = n_bca ? fetch_brick(X, nv, B0, C0, A0, tid) : 0.0;     // [b][c][a]
        const double r4 = need_x ? fetch_brick(X, nv, C0, A0, B0, tid) : 0.0;    // [c][a][b]
        const double r5 = n_cba ? fetch_brick(X, nv, C0, B0, A0, tid) : 0.0;     // [c][b][a]
        __syncthreads();
        store_brick(sW[0], r1, tid); store_brick(sW[1], r2, tid); store_brick(sW[2], r3, tid);
        store_brick(sW[3], r4, tid); store_brick(sW[4], r5, tid);
        __syncthreads();
        xacb = brick_at(sW[0], la, lc, lb);
        xbac = brick_at(sW[1], lb, la, lc);
        xbca = brick_at(sW[2], lb, lc, la);
        xcab = brick_at(sW[3], lc, la, lb);
        xcba = brick_at(sW[4], lc, lb, la);
    };
    double mabc, macb, mbac, mbca, mcab, mcba;
    gather(Wk, mabc, macb, mbac, mbca, mcab, mcba);
    double xn = 0.0;
    if (need_x) {
        double nabc, nacb, nbac, nbca, ncab, ncba;
        gather(Dd + (long long)kk * v3, nabc, nacb, nbac, nbca, ncab, ncba);
        xn = 8.0 * nabc - 4.0 * nbac - 4.0 * nacb - 4.0 * ncba + 2.0 * ncab + 2.0 * nbca;
    }
    if (a >= nv || b >= nv || c >= nv) return;
    const double d = (double)eps[i] + (double)eps[j] + (double)eps[k0 + kk] - (double)eps[no + a] - (double)eps[no + b] -
                     (double)eps[no + c];
    const double rd = 1.0 / d;
    mabc *= rd; macb *= rd; mbac *= rd; mbca *= rd; mcab *= rd; mcba *= rd; xn *= rd;
    const double xm = 8.0 * mabc - 4.0 * mbac - 4.0 * macb - 4.0 * mcba + 2.0 * mcab + 2.0 * mbca;
    const long long idx = (long long)kk * v3 + a * v2 + (long long)b * nv + c;
    if (o.M) o.M[idx] = (T)mabc;
    if (o.A1) o.A1[idx] = (T)(mabc - mcba);
    if (o.A2) o.A2[idx] = (T)(2.0 * mabc - macb - mcba);
    if (o.B1) o.B1[idx] = (T)(2.0 * (mabc - mbac));
    if (o.Z3) o.Z3[idx] = (T)(2.0 * (mabc - macb) - (mbac - mbca));
    if (o.P) o.P[idx] = (T)(2.0 * xm + xn);
    if (o.Q) o.Q[idx] = (T)(xm + xn);
}

template <typename T>
int launch_combine_brick(pccb_ctx* ctx, const void* w_dev, const void* dd_dev, int nk, int k0, int i, int j,
                         const void* eps_dev, int64_t no, int64_t nv, void* const* q) {
    const int nb = (int)((nv + TB_ - 1) / TB_);
    T3DensityOut<T> o{(T*)q[0], (T*)q[1], (T*)q[2], (T*)q[3], (T*)q[4], (T*)q[5], (T*)q[6]};
    dim3 grid((unsigned)(nb * nb * nb), (unsigned)nk);
    t3_combine_brick_kernel<T><<<grid, 512, 0, ctx->stream>>>((const T*)w_dev, (const T*)dd_dev, k0, i, j, (const T*)eps_dev,
                                                              (int)no, (int)nv, nb, o);
    PCCB_CHECK_LAUNCH(ctx, "t3_combine_brick_kernel");
    return PCCB_OK;
}

// PCCB_COMBINE_BRICK=0 restores the element-wise kernels (kept as the cross-check of the brick form)
bool use_combine_brick(int64_t nv) {
    static const bool on = [] { const char* e = getenv("PCCB_COMBINE_BRICK"); return !(e && e[0] == '0'); }();
    return on && nv <= 1290;        // nb^3 bricks in grid.x
}

template <typename T>
int triples_energy_t(pccb_ctx* ctx, const void* w_dev, const int32_t* ijk_dev, int ntrip, const void* t1,
                     const void* t2, const void* g, const void* fov, int64_t ld_fov, const void* eps, int64_t no,
                     int64_t nv, double* et_dev) {
    const int nb = (int)((nv + TB_ - 1) / TB_);
    const int npair = nb * (nb + 1) / 2;
    PCCB_REQUIRE(ctx, ntrip >= 1 && ntrip <= 65535, "pccb_triples_energy: ntrip=%d out of range", ntrip);
    PCCB_REQUIRE(ctx, nb <= 65535, "pccb_triples_energy: nv too large");
    int rc = pccb_ensure_partials(ctx, (size_t)ntrip * npair + 8);
    if (rc) return rc;
    double* partials = ctx->partial_dev;
    TriplesArgs<T> a;
    a.W = (const T*)w_dev; a.ijk = ijk_dev; a.t1 = (const T*)t1; a.t2 = (const T*)t2; a.g = (const T*)g;
    a.fov = (const T*)fov; a.eps = (const T*)eps; a.ld_fov = ld_fov; a.no = (int)no; a.nv = (int)nv; a.nb = nb;
    a.ntrip = ntrip; a.partials = partials;
    // non-canonical CTAs (TB > TA) exit at once, so only their partial slots stay unused
    dim3 grid(nb, nb, ntrip);
    triples_energy_kernel<T><<<grid, 512, 0, ctx->stream>>>(a);
    PCCB_CHECK_LAUNCH(ctx, "triples_energy_kernel");
    sum_rows_kernel<<<ntrip, 256, 0, ctx->stream>>>(partials, npair, et_dev);
    PCCB_CHECK_LAUNCH(ctx, "sum_rows_kernel");
    return PCCB_OK;
}

}  // namespace

extern "C" int pccb_triples_energy(pccb_ctx* ctx, int dtype, const void* w_dev, const int32_t* ijk_dev, int ntrip,
                                   const void* t1_dev, const void* t2_dev, const void* oovv_dev, const void* fov_dev,
                                   int64_t ld_fov, const void* eps_dev, int64_t no, int64_t nv, double* et_dev) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, w_dev && ijk_dev && t1_dev && t2_dev && oovv_dev && fov_dev && eps_dev && et_dev,
                 "pccb_triples_energy: null argument");
    PCCB_REQUIRE(ctx, no > 0 && nv > 0, "pccb_triples_energy: bad sizes");
    if (dtype == PCCB_F64)
        return triples_energy_t<double>(ctx, w_dev, ijk_dev, ntrip, t1_dev, t2_dev, oovv_dev, fov_dev, ld_fov,
                                        eps_dev, no, nv, et_dev);
    if (dtype == PCCB_F32)
        return triples_energy_t<float>(ctx, w_dev, ijk_dev, ntrip, t1_dev, t2_dev, oovv_dev, fov_dev, ld_fov,
                                       eps_dev, no, nv, et_dev);
    PCCB_FAIL(ctx, PCCB_ERR_ARG, "unknown dtype %d", dtype);
}

extern "C" int pccb_t3_combine(pccb_ctx* ctx, int dtype, const void* w_dev, int nk, int k0, int i, int j,
                               const void* eps_dev, int64_t no, int64_t nv, void* a1_dev, void* a2_dev) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, w_dev && eps_dev && a1_dev && a2_dev && nk >= 1 && no > 0 && nv > 0 && i >= 0 && j >= 0 &&
                          k0 >= 0 && k0 + nk <= no && i < no && j < no,
                 "pccb_t3_combine: bad arguments");
    const long long total = (long long)nk * nv * nv * nv;
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (dtype == PCCB_F64)
        t3_combine_kernel<double><<<(unsigned)blocks, 256, 0, ctx->stream>>>((const double*)w_dev, nk, k0, i, j,
                                                                             (const double*)eps_dev, (int)no, (int)nv,
                                                                             (double*)a1_dev, (double*)a2_dev);
    else if (dtype == PCCB_F32)
        t3_combine_kernel<float><<<(unsigned)blocks, 256, 0, ctx->stream>>>((const float*)w_dev, nk, k0, i, j,
                                                                            (const float*)eps_dev, (int)no, (int)nv,
                                                                            (float*)a1_dev, (float*)a2_dev);
    else
        PCCB_FAIL(ctx, PCCB_ERR_ARG, "unknown dtype %d", dtype);
    PCCB_CHECK_LAUNCH(ctx, "t3_combine_kernel");
    return PCCB_OK;
}

extern "C" int pccb_t3_density_combine(pccb_ctx* ctx, int dtype, const void* w_dev, const void* dd_dev, int nk, int k0, int i,
                                       int j, const void* eps_dev, int64_t no, int64_t nv, void* const* out7_dev_hostarray) {
    if (!ctx) return PCCB_ERR_ARG;
    PCCB_REQUIRE(ctx, w_dev && dd_dev && eps_dev && out7_dev_hostarray && nk >= 1 && no > 0 && nv > 0 && i >= 0 && j >= 0 &&
                          k0 >= 0 && k0 + nk <= no && i < no && j < no,
                 "pccb_t3_density_combine: bad arguments");
    void* const* q = out7_dev_hostarray;
    // all six permutations of W and Dd (P / Q requested: ccwfn.t3_density): the brick form, 1.36 -> 1.27 s at (16,120);
    // with one to three permutations (CC3: A1 / A2, M alone) the element-wise kernel is as fast or faster
    if ((q[5] || q[6]) && use_combine_brick(nv) && nk <= 65535 && (dtype == PCCB_F64 || dtype == PCCB_F32))
        return dtype == PCCB_F64 ? launch_combine_brick<double>(ctx, w_dev, dd_dev, nk, k0, i, j, eps_dev, no, nv, q)
                                 : launch_combine_brick<float>(ctx, w_dev, dd_dev, nk, k0, i, j, eps_dev, no, nv, q);
    const long long total = (long long)nk * nv * nv * nv;
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (dtype == PCCB_F64) {
        T3DensityOut<double> o{(double*)q[0], (double*)q[1], (double*)q[2], (double*)q[3], (double*)q[4], (double*)q[5], (double*)q[6]};
        t3_density_combine_kernel<double><<<(unsigned)blocks, 256, 0, ctx->stream>>>(
            (const double*)w_dev, (const double*)dd_dev, nk, k0, i, j, (const double*)eps_dev, (int)no, (int)nv, o);
    } else if (dtype == PCCB_F32) {
        T3DensityOut<float> o{(float*)q[0], (float*)q[1], (float*)q[2], (float*)q[3], (float*)q[4], (float*)q[5], (float*)q[6]};
        t3_density_combine_kernel<float><<<(unsigned)blocks, 256, 0, ctx->stream>>>(
            (const float*)w_dev, (const float*)dd_dev, nk, k0, i, j, (const float*)eps_dev, (int)no, (int)nv, o);
    } else {
        PCCB_FAIL(ctx, PCCB_ERR_ARG, "unknown dtype %d", dtype);
    }
    PCCB_CHECK_LAUNCH(ctx, "t3_density_combine_kernel");
    return PCCB_OK;
}
