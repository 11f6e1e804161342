// DTW stage downstream of the extractor, batched over (teacher, student) pairs of feature sequences
// (/root/reference/05_dtw_alignment_experiment/dtw_alignment.py):
//   dtw_distance_kernel    :532-552  composite frame distance 0.6 euclid - 0.25 (1 + cos) cos - 0.15 (1 + |corr|) corr
//   dtw_accumulate_kernel  :930-969  accumulated cost, D + min(1.2 up, 0.8 diagonal, 1.2 left), weighted borders
// float64 throughout, as numpy computes them.  The accumulated cost keeps the reference's operation order (no fused
// multiply-add), so it is bit-exact; the distance sums over the feature dimension in index order (numpy's pairwise /
// BLAS order differs in the last bits).
#pragma once
#include <cstdint>

namespace afx {

constexpr int kDtwTile = 32;
constexpr int kDtwDistThreads = 256;

struct DtwDistArgs {
  const double* x; const double* y; double* D;
  const long long* x_off; const long long* y_off; const long long* D_off;   // device [pairs], element offsets
  const int* n; const int* m;                                               // device [pairs]
  int dim, pair0;
};

// one CTA per 32 x 32 tile of one pair's matrix (blockIdx.y = pair, blockIdx.x strides over the tiles)
__global__ void __launch_bounds__(kDtwDistThreads) dtw_distance_kernel(DtwDistArgs a) {
  extern __shared__ __align__(16) double dsm[];
  const int dim = a.dim, ld = dim | 1;                       // odd leading dimension: conflict-free row reads
  double* xs = dsm;                                           // [32][ld]
  double* ys = xs + kDtwTile * ld;                            // [32][ld]
  double* st = ys + kDtwTile * ld;                            // [64][3]: mean, |v|, sqrt(sum centred^2)
  const int pair = a.pair0 + blockIdx.y;
  const int n = a.n[pair], m = a.m[pair];
  const double* __restrict__ X = a.x + a.x_off[pair];
  const double* __restrict__ Y = a.y + a.y_off[pair];
  double* __restrict__ D = a.D + a.D_off[pair];
  const int ti_n = (n + kDtwTile - 1) / kDtwTile, tj_n = (m + kDtwTile - 1) / kDtwTile;
  const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
  for (int tile = blockIdx.x; tile < ti_n * tj_n; tile += gridDim.x) {
    const int i0 = (tile / tj_n) * kDtwTile, j0 = (tile % tj_n) * kDtwTile;
    __syncthreads();
    for (int e = tid; e < kDtwTile * dim; e += kDtwDistThreads) {
      const int r = e / dim, k = e - r * dim;
      xs[r * ld + k] = i0 + r < n ? X[(size_t)(i0 + r) * dim + k] : 0.0;
      ys[r * ld + k] = j0 + r < m ? Y[(size_t)(j0 + r) * dim + k] : 0.0;
    }
    __syncthreads();
    if (tid < 2 * kDtwTile) {                                // per-row statistics
      const double* v = (tid < kDtwTile ? xs : ys) + (tid & (kDtwTile - 1)) * ld;
      double s = 0.0, q = 0.0;
      for (int k = 0; k < dim; ++k) { s += v[k]; q = fma(v[k], v[k], q); }
      const double mean = s / (double)dim;
      double c = 0.0;
      for (int k = 0; k < dim; ++k) { const double d = v[k] - mean; c = fma(d, d, c); }
      st[tid * 3 + 0] = mean; st[tid * 3 + 1] = sqrt(q); st[tid * 3 + 2] = sqrt(c);
    }
    __syncthreads();
    const double* yv = ys + tx * ld;
    const double my = st[(kDtwTile + tx) * 3], ny = st[(kDtwTile + tx) * 3 + 1], sy = st[(kDtwTile + tx) * 3 + 2];
#pragma unroll
    for (int u = 0; u < kDtwTile / 8; ++u) {
      const int r = ty + 8 * u;
      const double* xv = xs + r * ld;
      const double mx = st[r * 3], nx = st[r * 3 + 1], sx = st[r * 3 + 2];
      double e2 = 0.0, dot = 0.0, cov = 0.0;
      for (int k = 0; k < dim; ++k) {
        const double p = xv[k], q = yv[k], d = p - q;
        e2 = fma(d, d, e2);
        dot = fma(p, q, dot);
        cov = fma(p - mx, q - my, cov);
      }
      const double cosv = dot / (nx * ny + 1e-8);
      double corr = cov / sx / sy;                            // np.corrcoef: one division per standard deviation
      if (corr > 1.0) corr = 1.0;                             // np.clip keeps NaN (constant frame), as these do
      if (corr < -1.0) corr = -1.0;
      const double w2 = 0.25 * (1.0 + cosv), w3 = 0.15 * (1.0 + fabs(corr));
      const double val = 0.6 * sqrt(e2) - w2 * cosv - w3 * corr;
      if (i0 + r < n && j0 + tx < m) D[(size_t)(i0 + r) * m + j0 + tx] = val;
    }
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

// one CTA per pair; strips of THREADS rows, thread r owns row s0 + r and walks the anti-diagonals: at step d it is at
// column d - r + 1, takes `up` from the thread above (shared memory, written one step earlier), keeps `diag` and
// `left` in registers.  One barrier per anti-diagonal.
template <int THREADS>
__global__ void __launch_bounds__(THREADS) dtw_accumulate_kernel(DtwAccArgs a) {
  __shared__ double sh[2][THREADS];
  const int pair = a.pair0 + blockIdx.x;
  const int n = a.n[pair], m = a.m[pair];
  if (n <= 0 || m <= 0) return;
  const double* __restrict__ D = a.D + a.D_off[pair];
  double* A = a.A + a.D_off[pair];
  const int tid = threadIdx.x;
  // borders (:936-945): serial prefix sums with the 1 + 0.1 k weight; column 0 by warp 0, row 0 by warp 1
  if (tid == 0) {
    double acc = D[0];
    A[0] = acc;
    for (int i = 1; i < n; ++i) {
      acc = __dadd_rn(acc, __dmul_rn(D[(size_t)i * m], __dadd_rn(1.0, __dmul_rn(0.1, (double)i))));
      A[(size_t)i * m] = acc;
    }
  } else if (tid == 32) {
    double acc = D[0];
    for (int j = 1; j < m; ++j) {
      acc = __dadd_rn(acc, __dmul_rn(D[j], __dadd_rn(1.0, __dmul_rn(0.1, (double)j))));
      A[j] = acc;
    }
  }
  __syncthreads();
  for (int s0 = 1; s0 < n; s0 += THREADS) {
    const int rows = min(THREADS, n - s0);
    const int i = s0 + tid;
    const bool mine = tid < rows;
    const double* Drow = D + (size_t)i * m;
    double* Arow = A + (size_t)i * m;
    const double* Aup = A + (size_t)(i - 1) * m;              // read by thread 0 only: the finished row above the strip
    double left = mine ? Arow[0] : 0.0;
    double diag = mine ? Aup[0] : 0.0;
    const int steps = (m - 1) + rows - 1;
    double d_next = (mine && tid == 0 && m > 1) ? Drow[1] : 0.0;
    double up_next = (mine && tid == 0 && m > 1) ? Aup[1] : 0.0;
    for (int d = 0; d < steps; ++d) {
      const int j = d - tid + 1;
      const bool active = mine && j >= 1 && j < m;
      const double dv = d_next, upg = up_next;
      // prefetch the next step's operands off the dependency chain
      const int jn = j + 1;
      if (mine && jn >= 1 && jn < m) { d_next = Drow[jn]; if (tid == 0) up_next = Aup[jn]; }
      if (active) {
        const double up = tid == 0 ? upg : sh[(d + 1) & 1][tid - 1];
        const double v = dtw_step(dv, up, diag, left);
        Arow[j] = v;
        sh[d & 1][tid] = v;
        left = v; diag = up;
      }
      __syncthreads();
    }
    __threadfence_block();
    __syncthreads();   // the strip's last row is read from global memory by the next strip
  }
}

}  // namespace afx
