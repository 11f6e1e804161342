// DTW stage downstream of the extractor, batched over (teacher, student) pairs of feature sequences
// (/root/reference/05_dtw_alignment_experiment/dtw_alignment.py):
//   dtw_distance_kernel    :532-552  composite frame distance 0.6 euclid - 0.25 (1 + cos) cos - 0.15 (1 + |corr|) corr
//                                    (one float64 tensor-core product per tile + a per-cell epilogue)
//   dtw_accumulate_kernel  :930-969  accumulated cost, D + min(1.2 up, 0.8 diagonal, 1.2 left), weighted borders
// float64 throughout, as numpy computes them.  The accumulated cost keeps the reference's operation order (no fused
// multiply-add), so it is bit-exact; the distance sums over the feature dimension in index order (numpy's pairwise /
// BLAS order differs in the last bits).
#pragma once
#include <cstdint>

namespace afx {

constexpr int kDtwTile = 64;            // output tile of one CTA: 64 x 64 cells, one 32 x 32 quadrant per warp
constexpr int kDtwDistThreads = 128;

struct DtwDistArgs {
  const double* x; const double* y; double* D;
  const long long* x_off; const long long* y_off; const long long* D_off;   // device [pairs], element offsets
  const int* n; const int* m;                                               // device [pairs]
  int dim, ld, pair0;                                                       // ld: shared-memory row pitch (doubles), ld % 8 == 4
};

// D = A * B + C on one 8 x 8 x 4 float64 tensor-core tile (DMMA): lane l holds A[l / 4][l % 4], B[l % 4][l / 4] and
// C[l / 4][2 (l % 4) + {0, 1}]
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// per-cell arithmetic of the distance on one lane's two accumulators of an 8 x 8 DMMA tile (kept out of line: 16 inlined
// copies of the float64 sqrt / reciprocal sequences overflow the instruction cache)
__device__ __noinline__ void dtw_distance_epilogue(double cov0, double cov1, const double* st, int r, int cl, int i0, int j0,
                                                   int n, int m, int dim, const double* __restrict__ X,
                                                   const double* __restrict__ Y, double* __restrict__ D) {
  const double mx = st[r * 4], qx = st[r * 4 + 1], nx = st[r * 4 + 2], rsx = st[r * 4 + 3];
  double out[2];
#pragma unroll
  for (int e = 0; e < 2; ++e) {
    const int c = cl + e;
    const double* sy = st + (kDtwTile + c) * 4;
    const double cov = e ? cov1 : cov0;
    const double dot = fma((double)dim * mx, sy[0], cov);
    double e2 = fma(-2.0, dot, qx + sy[1]);
    if (e2 < 1e-2 * (qx + sy[1]) && i0 + r < n && j0 + c < m) {     // cancellation: sum the differences themselves
      const double* xr = X + (size_t)(i0 + r) * dim;
      const double* yr = Y + (size_t)(j0 + c) * dim;
      e2 = 0.0;
#pragma unroll 1
      for (int k = 0; k < dim; ++k) { const double d = xr[k] - yr[k]; e2 = fma(d, d, e2); }
    }
    const double cosv = dot * __drcp_rn(fma(nx, sy[2], 1e-8));
    double corr = cov * rsx * sy[3];                    // np.corrcoef; 0 * inf = NaN for a constant frame, as 0 / 0 there
    if (corr > 1.0) corr = 1.0;                         // np.clip keeps NaN, as these do
    if (corr < -1.0) corr = -1.0;
    const double w2 = 0.25 * (1.0 + cosv), w3 = 0.15 * (1.0 + fabs(corr));
    out[e] = 0.6 * sqrt(fmax(e2, 0.0)) - w2 * cosv - w3 * corr;
  }
  const int c0 = j0 + cl;
  if (i0 + r < n) {
    double* dst = D + (size_t)(i0 + r) * m + c0;
    if (c0 < m) dst[0] = out[0];
    if (c0 + 1 < m) dst[1] = out[1];
  }
}

// One CTA per 64 x 64 tile of one pair's matrix (blockIdx.y = pair, blockIdx.x strides over the tiles).  The three sums
// over the feature dimension collapse into ONE float64 matrix product on the tensor cores, cov = Xc Yc^T of the
// mean-centred rows: dot = cov + dim mean_x mean_y, |x - y|^2 = |x|^2 + |y|^2 - 2 dot.  Where that difference cancels
// (frames closer than 10 % of their norms: near-identical teacher / student frames) the cell recomputes sum (x - y)^2
// directly from global memory, so an exactly equal pair of frames still gives exactly 0.
__global__ void __launch_bounds__(kDtwDistThreads) dtw_distance_kernel(DtwDistArgs a) {
  extern __shared__ __align__(16) double dsm[];
  const int dim = a.dim, ld = a.ld;
  double* xs = dsm;                                           // [64][ld] x rows (centred in place)
  double* ys = xs + kDtwTile * ld;                            // [64][ld] y rows
  double* st = ys + kDtwTile * ld;                            // [128][4]: mean, |v|^2, |v|, 1 / sqrt(sum centred^2)
  const int pair = a.pair0 + blockIdx.y;
  const int n = a.n[pair], m = a.m[pair];
  const double* __restrict__ X = a.x + a.x_off[pair];
  const double* __restrict__ Y = a.y + a.y_off[pair];
  double* __restrict__ D = a.D + a.D_off[pair];
  const int ti_n = (n + kDtwTile - 1) / kDtwTile, tj_n = (m + kDtwTile - 1) / kDtwTile;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ksteps = (dim + 3) >> 2;
  for (int tile = blockIdx.x; tile < ti_n * tj_n; tile += gridDim.x) {
    const int i0 = (tile / tj_n) * kDtwTile, j0 = (tile % tj_n) * kDtwTile;
    __syncthreads();
    for (int e = tid; e < kDtwTile * ld; e += kDtwDistThreads) {
      const int r = e / ld, k = e - r * ld;
      xs[e] = (k < dim && i0 + r < n) ? X[(size_t)(i0 + r) * dim + k] : 0.0;
      ys[e] = (k < dim && j0 + r < m) ? Y[(size_t)(j0 + r) * dim + k] : 0.0;
    }
    __syncthreads();
    {                                                         // per-row statistics, rows centred in place
      double* v = (tid < kDtwTile ? xs : ys) + (tid & (kDtwTile - 1)) * ld;
      // the row is zero beyond dim (ld is a multiple of 4): four independent chains per sum
      double s4[4] = {0, 0, 0, 0}, q4[4] = {0, 0, 0, 0}, c4[4] = {0, 0, 0, 0};
      const int k4 = (dim + 3) & ~3;
      // each thread starts at its own column (rows are ld = 4 mod 8 doubles apart: the same column would be a 4-way
      // bank conflict)
      const int rot = (tid & 15) < k4 ? (tid & 15) : 0;
      for (int kk = 0; kk < k4; kk += 4)
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          int k = kk + u + rot; k -= k >= k4 ? k4 : 0;
          const double x = v[k]; s4[u] += x; q4[u] = fma(x, x, q4[u]);
        }
      const double mean = ((s4[0] + s4[1]) + (s4[2] + s4[3])) / (double)dim;
      const double q = (q4[0] + q4[1]) + (q4[2] + q4[3]);
      for (int kk = 0; kk < k4; kk += 4)
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          int k = kk + u + rot; k -= k >= k4 ? k4 : 0;
          const double d = k < dim ? v[k] - mean : 0.0;
          v[k] = d; c4[u] = fma(d, d, c4[u]);
        }
      const double c = (c4[0] + c4[1]) + (c4[2] + c4[3]);
      st[tid * 4 + 0] = mean; st[tid * 4 + 1] = q; st[tid * 4 + 2] = sqrt(q); st[tid * 4 + 3] = 1.0 / sqrt(c);
    }
    __syncthreads();
    // ---- cov = Xc Yc^T for this warp's 32 x 32 quadrant: 4 x 4 DMMA tiles, k in steps of 4
    const int wi = (warp >> 1) * 32, wj = (warp & 1) * 32;
    double acc[4][4][2];
#pragma unroll
    for (int mb = 0; mb < 4; ++mb)
#pragma unroll
      for (int nb = 0; nb < 4; ++nb) acc[mb][nb][0] = acc[mb][nb][1] = 0.0;
    const double* xa = xs + (wi + (lane >> 2)) * ld + (lane & 3);
    const double* yb = ys + (wj + (lane >> 2)) * ld + (lane & 3);
    for (int ks = 0; ks < ksteps; ++ks) {
      double af[4], bf[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) { af[q] = xa[q * 8 * ld + 4 * ks]; bf[q] = yb[q * 8 * ld + 4 * ks]; }
#pragma unroll
      for (int mb = 0; mb < 4; ++mb)
#pragma unroll
        for (int nb = 0; nb < 4; ++nb) dmma884(acc[mb][nb][0], acc[mb][nb][1], af[mb], bf[nb]);
    }
    // ---- epilogue: lane holds rows wi + 8 mb + lane / 4, columns wj + 8 nb + 2 (lane % 4) + {0, 1}
#pragma unroll
    for (int mb = 0; mb < 4; ++mb)
#pragma unroll
      for (int nb = 0; nb < 4; ++nb)
        dtw_distance_epilogue(acc[mb][nb][0], acc[mb][nb][1], st, wi + 8 * mb + (lane >> 2), wj + 8 * nb + 2 * (lane & 3),
                              i0, j0, n, m, dim, X, Y, D);
  }
}

struct DtwAccArgs {
  const double* D; double* A;
  const long long* D_off;
  const int* n; const int* m;
  int pair0;
};

__device__ __forceinline__ double dtw_step(double d, double up, double diag, double left) {
  // min(vertical * 1.2, diagonal * 0.8, horizontal * 1.2) with Python's min: the first of equal / unordered values stays
  double mn = __dmul_rn(up, 1.2);
  const double dg = __dmul_rn(diag, 0.8), hz = __dmul_rn(left, 1.2);
  if (dg < mn) mn = dg;
  if (hz < mn) mn = hz;
  return __dadd_rn(d, mn);
}

// One WARP per pair.  Row 0 is a serial prefix (:942-945).  The rest of the matrix is swept in bands of 32 rows; lane r
// owns row i0 + r and at step t sits at column t - r + 1 (an anti-diagonal of the band), takes `up` from lane r - 1 by
// shuffle (the value that lane produced one step earlier) and keeps `diag` and `left` in registers: no barrier on the
// dependency chain, and only the interior recurrence inside the loop (column 0, :936-940, is a 32-step prefix over
// the lanes at the head of each band).  D enters and A leaves through a circular buffer of 96 columns per row in shared
// memory: cp.async fills the 32-column block after the one lane 0 is in, each cell is overwritten in place by its
// accumulated cost and finished blocks are written back row by row (256-byte rows both ways, where a lane-per-row walk of
// global memory would touch one sector per lane per step).  Row 0 of the buffer holds the boundary row A[i0 - 1].
constexpr int kDtwAccWarps = 4;
constexpr int kDtwRing = 96;                                  // columns in the circular buffer: blocks cb - 1, cb, cb + 1
constexpr int kDtwLd = 98;                                    // doubles per buffer row: lane r at column c - r -> 97 r + c, distinct bank pairs
constexpr int kDtwRingDoubles = 33 * kDtwLd;
constexpr int kDtwAccSmem = kDtwAccWarps * kDtwRingDoubles * (int)sizeof(double);

__device__ __forceinline__ void cp_async8(double* dst_smem, const double* src) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(dst_smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(d), "l"(src) : "memory");
}

__device__ __forceinline__ double dtw_border_weight(int k) { return __dadd_rn(1.0, __dmul_rn(0.1, (double)k)); }

__global__ void __launch_bounds__(kDtwAccWarps * 32) dtw_accumulate_kernel(DtwAccArgs a, int n_pairs) {
  extern __shared__ __align__(16) double dtw_sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int pair = a.pair0 + blockIdx.x * kDtwAccWarps + warp;
  if (pair >= n_pairs) return;
  const int n = a.n[pair], m = a.m[pair];
  if (n <= 0 || m <= 0) return;
  const double* __restrict__ D = a.D + a.D_off[pair];
  double* A = a.A + a.D_off[pair];
  double* ring = dtw_sm + (size_t)warp * kDtwRingDoubles;
  const int nb = (m + 31) >> 5;
  // ---- row 0: A[0][j] = A[0][j - 1] + D[0][j] (1 + 0.1 j), additions strictly in column order
  double carry = 0.0;
  for (int j0 = 0; j0 < m; j0 += 32) {
    const int j = j0 + lane;
    double term = 0.0;
    if (j < m) { const double d = D[j]; term = j == 0 ? d : __dmul_rn(d, dtw_border_weight(j)); }
    double mine = 0.0;
#pragma unroll 8
    for (int l = 0; l < 32; ++l) {
      const double tl = __shfl_sync(0xffffffffu, term, l);
      carry = (j0 + l == 0) ? tl : __dadd_rn(carry, tl);
      if (lane == l) mine = carry;
    }
    if (j < m) A[j] = mine;
  }
  double col0 = __shfl_sync(0xffffffffu, D[0], 0);            // A[i0 - 1][0] of the first band = A[0][0] = D[0][0]
  __threadfence_block();
  __syncwarp();
  for (int i0 = 1; i0 < n; i0 += 32) {
    const int i = i0 + lane;
    const bool row_ok = i < n;
    const int rows = min(32, n - i0);
    // fill of block cb: buffer rows 1..rows from D, row 0 from the finished row above the band
    auto fill = [&](int cb, int rc) {
      const int j = (cb << 5) + lane;
      if (cb < nb && j < m) {
        double* dst = ring + rc + lane;
        cp_async8(dst, A + (size_t)(i0 - 1) * m + j);
        const double* src = D + (size_t)i0 * m + j;
        for (int rr = 0; rr < rows; ++rr) { dst += kDtwLd; cp_async8(dst, src); src += m; }
      }
      asm volatile("cp.async.commit_group;\n" ::: "memory");
    };
    fill(0, 0);
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    __syncwarp();
    // ---- column 0 of the band: A[i][0] = A[i - 1][0] + D[i][0] (1 + 0.1 i), serial over the lanes
    double left, diag;
    {
      const double term = row_ok ? __dmul_rn(ring[(lane + 1) * kDtwLd], dtw_border_weight(i)) : 0.0;
      double mine = 0.0, acc = col0;
#pragma unroll 8
      for (int l = 0; l < 32; ++l) {
        acc = __dadd_rn(acc, __shfl_sync(0xffffffffu, term, l));
        if (lane == l) mine = acc;
      }
      const double above = __shfl_up_sync(0xffffffffu, mine, 1);
      left = mine;
      diag = lane == 0 ? col0 : above;
      if (row_ok) ring[(lane + 1) * kDtwLd] = mine;
      col0 = __shfl_sync(0xffffffffu, mine, rows - 1);        // boundary of the next band
    }
    __syncwarp();
    double v = left;
    int j = -lane;                                            // lane r's column at the step where lane 0 is at column 0
    int jm = (kDtwRing - lane) % kDtwRing;                    // j mod 96
    double* cell = ring + (lane + 1) * kDtwLd;                // this lane's buffer row
    int rc_fill = 32, rc_back = 0;                            // buffer columns of block cb + 1 / block cb - 2
    for (int cb = 0; cb < nb + 2; ++cb) {
      __syncwarp();
      if (cb >= 2) {                                          // block cb - 2 is complete: write it back
        const int jb = ((cb - 2) << 5) + lane;
        if (jb < m) {
          const double* src = ring + kDtwLd + rc_back + lane;
          double* dst = A + (size_t)i0 * m + jb;
          for (int rr = 0; rr < rows; ++rr) { *dst = *src; src += kDtwLd; dst += m; }
        }
        rc_back = rc_back == 64 ? 0 : rc_back + 32;
      }
      __syncwarp();
      fill(cb + 1, rc_fill);                                  // into the columns just written back
      rc_fill = rc_fill == 64 ? 0 : rc_fill + 32;
      asm volatile("cp.async.wait_group 1;\n" ::: "memory");  // block cb has landed
      __syncwarp();
      if (cb >= nb + 1) continue;                             // only write-backs left
#pragma unroll 4
      for (int s = 0; s < 32; ++s) {
        const double vup = __shfl_up_sync(0xffffffffu, v, 1);
        const bool active = row_ok && j >= 1 && j < m;
        const double d = cell[jm];
        const double up = lane == 0 ? ring[jm] : vup;
        // min(up * 1.2, diag * 0.8, left * 1.2), first of equal / unordered values kept (Python's min): the two comparisons
        // against up * 1.2 are independent, the third does not wait for the shuffle
        const double u12 = __dmul_rn(up, 1.2), dg = __dmul_rn(diag, 0.8), hz = __dmul_rn(left, 1.2);
        const double p = hz < dg ? hz : dg;
        const double q = hz < u12 ? hz : u12;
        const double nv = __dadd_rn(d, dg < u12 ? p : q);
        if (active) { cell[jm] = nv; v = nv; left = nv; diag = up; }
        ++j;
        jm = jm == kDtwRing - 1 ? 0 : jm + 1;
      }
    }
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    __threadfence_block();
    __syncwarp();   // the band's last row is the next band's boundary row
  }
}

}  // namespace afx
