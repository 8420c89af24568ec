// Moments and projection for the Memory Galaxy PCA.
//
// The reference reduces the embeddings of up to `Cap` entities to PCADims columns before t-SNE
// (ee/pkg/memory/projection/pca.go:28-67, project.go:120-136): gonum centres the columns, takes an SVD
// and multiplies the UNCENTRED matrix by the first k right singular vectors, all in float64.  The work
// that grows with the number of rows is done here from the fp32 master rows already resident in HBM:
//
//   pca_colsum_kernel   column sums of the selected rows              n * D       fp64 adds
//   pca_cov_kernel      upper-triangle tiles of sum (x-mu)(x-mu)^T    n * D^2     DFMA (the dominant kernel)
//   pca_reduce_kernel   fixed-order sum of the row-split partials, division by n-1, mirror to the lower half
//   pca_project_kernel  X * V for a D x k basis                       2 * n * D * k
//
// The D x D symmetric eigenproblem between the two (independent of n) is solved on the host.  Everything
// is fp64 like the reference; partial sums are combined in a fixed order, so a given input always gives
// the same bits.  Roofline: fp64 pipe (B200 nominal 40 TFLOP/s; not in MEASURED_PEAKS.json).
#include <algorithm>

#include "kernels.cuh"

namespace omr {

namespace {

constexpr int kCovTile = 128;  // output tile edge
constexpr int kCovStep = 8;    // rows staged per step

// grid.x = number of row chunks; partial is [gridDim.x][dim]
__global__ void __launch_bounds__(256) pca_colsum_kernel(ShardView sh, const int64_t* __restrict__ rows, int64_t n,
                                                         int64_t chunk, double* __restrict__ partial) {
  const int64_t r0 = (int64_t)blockIdx.x * chunk;
  const int64_t r1 = min(n, r0 + chunk);
  for (int c = threadIdx.x; c < sh.dim; c += blockDim.x) {
    double acc = 0.0;
    for (int64_t r = r0; r < r1; r++) acc += (double)sh.rows_f32[rows[r] * sh.pitch + c];
    partial[(int64_t)blockIdx.x * sh.dim + c] = acc;
  }
}

__global__ void __launch_bounds__(256) pca_mean_kernel(const double* __restrict__ partial, int n_partial, int dim,
                                                       int64_t n, double* __restrict__ mean) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= dim) return;
  double acc = 0.0;
  for (int p = 0; p < n_partial; p++) acc += partial[(int64_t)p * dim + c];
  mean[c] = acc / (double)n;
}

// grid = (upper-triangle tile index, row split).  Thread (ty, tx) of the 16 x 16 layout owns the 8 x 8
// outputs (ty + 16 a, tx + 16 b): operands of one k-step are then read from shared memory at consecutive
// addresses across a half-warp, without bank conflicts.
__global__ void __launch_bounds__(256, 1) pca_cov_kernel(ShardView sh, const int64_t* __restrict__ rows, int64_t n,
                                                         int64_t chunk, const double* __restrict__ mean,
                                                         int n_tiles_edge, double* __restrict__ partial) {
  __shared__ double As[kCovStep][kCovTile];
  __shared__ double Bs[kCovStep][kCovTile];
  // decode the upper-triangle tile (ti <= tj) from the linear index
  int ti = 0, rem = blockIdx.x;
  while (rem >= n_tiles_edge - ti) {
    rem -= n_tiles_edge - ti;
    ti++;
  }
  const int tj = ti + rem;
  const int i0 = ti * kCovTile, j0 = tj * kCovTile;
  const int64_t r0 = (int64_t)blockIdx.y * chunk;
  const int64_t r1 = min(n, r0 + chunk);
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  // staging role: one column of each tile, rows lr, lr + 2, lr + 4, lr + 6 of the step
  const int lc = tid & (kCovTile - 1), lr = tid >> 7;
  const bool a_ok = i0 + lc < sh.dim, b_ok = j0 + lc < sh.dim;
  const double mu_a = a_ok ? mean[i0 + lc] : 0.0;
  const double mu_b = b_ok ? mean[j0 + lc] : 0.0;

  double acc[8][8];
#pragma unroll
  for (int a = 0; a < 8; a++)
#pragma unroll
    for (int b = 0; b < 8; b++) acc[a][b] = 0.0;

  for (int64_t rb = r0; rb < r1; rb += kCovStep) {
    float fa[4], fb[4];
    bool live[4];
#pragma unroll
    for (int s = 0; s < 4; s++) {
      const int64_t r = rb + lr + 2 * s;
      live[s] = r < r1;
      const float* src = live[s] ? sh.rows_f32 + rows[r] * sh.pitch : nullptr;
      fa[s] = (live[s] && a_ok) ? src[i0 + lc] : 0.0f;
      fb[s] = (live[s] && b_ok) ? src[j0 + lc] : 0.0f;
    }
    __syncthreads();  // the previous step's operands have been consumed
#pragma unroll
    for (int s = 0; s < 4; s++) {
      As[lr + 2 * s][lc] = (live[s] && a_ok) ? (double)fa[s] - mu_a : 0.0;
      Bs[lr + 2 * s][lc] = (live[s] && b_ok) ? (double)fb[s] - mu_b : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < kCovStep; kk++) {
      double av[8], bv[8];
#pragma unroll
      for (int a = 0; a < 8; a++) av[a] = As[kk][ty + 16 * a];
#pragma unroll
      for (int b = 0; b < 8; b++) bv[b] = Bs[kk][tx + 16 * b];
#pragma unroll
      for (int a = 0; a < 8; a++)
#pragma unroll
        for (int b = 0; b < 8; b++) acc[a][b] = fma(av[a], bv[b], acc[a][b]);
    }
  }
  double* out = partial + (int64_t)blockIdx.y * sh.dim * sh.dim;
#pragma unroll
  for (int a = 0; a < 8; a++) {
    const int i = i0 + ty + 16 * a;
    if (i >= sh.dim) continue;
#pragma unroll
    for (int b = 0; b < 8; b++) {
      const int j = j0 + tx + 16 * b;
      if (j < sh.dim) out[(int64_t)i * sh.dim + j] = acc[a][b];
    }
  }
}

// cov[i][j] for i <= j in computed tiles; the lower half is the mirror image.
__global__ void __launch_bounds__(256) pca_reduce_kernel(const double* __restrict__ partial, int n_split, int dim,
                                                         double inv_div, double* __restrict__ cov) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= (int64_t)dim * dim) return;
  const int i = (int)(e / dim), j = (int)(e % dim);
  // element (i, j) was written iff its tile is on or above the diagonal
  const bool upper = (i / kCovTile) <= (j / kCovTile);
  const int64_t src = upper ? (int64_t)i * dim + j : (int64_t)j * dim + i;
  double acc = 0.0;
  for (int s = 0; s < n_split; s++) acc += partial[(int64_t)s * dim * dim + src];
  cov[e] = acc * inv_div;
}

// grid = (ceil(n / 16), ceil(k / 64)).  Thread (r, cl) owns out[row r][cl + 16 m], m < 4.
__global__ void __launch_bounds__(256) pca_project_kernel(ShardView sh, const int64_t* __restrict__ rows, int64_t n,
                                                          const double* __restrict__ basis, int k,
                                                          double* __restrict__ out) {
  __shared__ double xs[16][33];
  __shared__ double vs[32][64];
  const int tid = threadIdx.x;
  const int r = tid >> 4, cl = tid & 15;
  const int64_t row0 = (int64_t)blockIdx.x * 16;
  const int c0 = blockIdx.y * 64;
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  for (int j0 = 0; j0 < sh.dim; j0 += 32) {
    __syncthreads();
    for (int e = tid; e < 16 * 32; e += 256) {
      const int rr = e >> 5, jj = e & 31;
      const int64_t gr = row0 + rr;
      xs[rr][jj] = (gr < n && j0 + jj < sh.dim) ? (double)sh.rows_f32[rows[gr] * sh.pitch + j0 + jj] : 0.0;
    }
    for (int e = tid; e < 32 * 64; e += 256) {
      const int jj = e >> 6, cc = e & 63;
      vs[jj][cc] = (j0 + jj < sh.dim && c0 + cc < k) ? basis[(int64_t)(j0 + jj) * k + c0 + cc] : 0.0;
    }
    __syncthreads();
#pragma unroll 8
    for (int jj = 0; jj < 32; jj++) {
      const double x = xs[r][jj];
#pragma unroll
      for (int m = 0; m < 4; m++) acc[m] = fma(x, vs[jj][cl + 16 * m], acc[m]);
    }
  }
  const int64_t gr = row0 + r;
  if (gr < n) {
#pragma unroll
    for (int m = 0; m < 4; m++) {
      const int c = c0 + cl + 16 * m;
      if (c < k) out[gr * k + c] = acc[m];
    }
  }
}

}  // namespace

int pca_mean_chunks(int64_t n) { return (int)std::max<int64_t>(1, std::min<int64_t>(592, n / 8)); }

int pca_cov_splits(int64_t n, int dim, int num_sms) {
  const int edge = (dim + kCovTile - 1) / kCovTile;
  const int tiles = edge * (edge + 1) / 2;
  int splits = (2 * num_sms + tiles - 1) / tiles;
  const int64_t by_rows = std::max<int64_t>(1, n / 64);
  splits = (int)std::min<int64_t>(std::min(splits, 64), by_rows);
  return std::max(1, splits);
}

int launch_pca_moments(const ShardView& sh, const int64_t* rows, int64_t n, double* colsum_partial, int n_chunks,
                       double* mean, double* cov_partial, int n_splits, double* cov, cudaStream_t st) {
  const int64_t chunk = (n + n_chunks - 1) / n_chunks;
  pca_colsum_kernel<<<n_chunks, 256, 0, st>>>(sh, rows, n, chunk, colsum_partial);
  OMR_CUDA_OK(cudaGetLastError());
  pca_mean_kernel<<<(sh.dim + 255) / 256, 256, 0, st>>>(colsum_partial, n_chunks, sh.dim, n, mean);
  OMR_CUDA_OK(cudaGetLastError());
  const int edge = (sh.dim + kCovTile - 1) / kCovTile;
  const int tiles = edge * (edge + 1) / 2;
  int64_t cchunk = (n + n_splits - 1) / n_splits;
  cchunk = (cchunk + kCovStep - 1) / kCovStep * kCovStep;
  pca_cov_kernel<<<dim3((unsigned)tiles, (unsigned)n_splits), 256, 0, st>>>(sh, rows, n, cchunk, mean, edge, cov_partial);
  OMR_CUDA_OK(cudaGetLastError());
  const int64_t elems = (int64_t)sh.dim * sh.dim;
  // gonum's PrincipalComponents scales by 1 / (n - 1) (sample covariance); n >= 2 is the caller's guard
  pca_reduce_kernel<<<(unsigned)((elems + 255) / 256), 256, 0, st>>>(cov_partial, n_splits, sh.dim,
                                                                      1.0 / (double)(n - 1), cov);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_pca_project(const ShardView& sh, const int64_t* rows, int64_t n, const double* basis, int k, double* out,
                       cudaStream_t st) {
  pca_project_kernel<<<dim3((unsigned)((n + 15) / 16), (unsigned)((k + 63) / 64)), 256, 0, st>>>(sh, rows, n, basis, k,
                                                                                                  out);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace omr
