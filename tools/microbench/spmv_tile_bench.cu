// Microbenchmark: the column-blocked shared-memory-gather SpMV (spmv_tile.cuh) on synthetic
// matrices of BASELINE config 2's shapes, next to the numbers of spmv_sell_bench.cu:
//   A-like : m = 1e6 rows x n = 1e5 columns, exactly 10 random columns per row  (4 column blocks)
//   At-like: n = 1e5 rows x m = 1e6 columns, 100 +- 10 random columns per row   (37 column blocks)
// Tiles are built on the host here (the product builds them on the device, dev_mats.cu).
// Prints the tile kernel time, the combine (row sum + rho scaling) kernel time and the check
// against a host CSR product.
#include "spmv_tile_proto.cuh"
#include <algorithm>
#include <cmath>
#include <vector>

static unsigned long long rs = 88172645463325252ull;
static inline unsigned long long rnd() { rs ^= rs << 13; rs ^= rs >> 7; rs ^= rs << 17; return rs; }
struct Host { int rows, cols; std::vector<int> rp, col; std::vector<double> val; };
template <class T> T *up(const std::vector<T> &h, size_t extra = 64) {
  T *d; cudaMalloc(&d, (h.size() + extra) * sizeof(T)); cudaMemset(d, 0, (h.size() + extra) * sizeof(T));
  cudaMemcpy(d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice); return d;
}
template <class F> float timeit(F f, int reps = 20) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int r = 0; r < 3; ++r) f();
  cudaDeviceSynchronize();
  cudaEventRecord(a);
  for (int r = 0; r < reps; ++r) f();
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms / reps * 1000.f;
}

// One CTA per window of 1024 rows: the window's partial sums of a group of blocks are staged in
// shared memory (coalesced), every row picks its own through inv[c][row]; sums in block order.
#define COMB_G 4
__global__ void __launch_bounds__(1024) k_combine(const double *partial, const unsigned short *inv, int nblocks, int rows,
                                                   int rows_pad, const double *rho, double *w) {
  __shared__ double sp[COMB_G][TILE_WIN];
  const int tid = threadIdx.x, win = blockIdx.x;
  const int i = win * TILE_WIN + tid;
  const double rh = (i < rows) ? rho[i] : 0.0;
  double acc = 0.0;
  for (int c0 = 0; c0 < nblocks; c0 += COMB_G) {
    const int g = min(COMB_G, nblocks - c0);
    unsigned short q[COMB_G]; double v[COMB_G];
#pragma unroll
    for (int k = 0; k < COMB_G; ++k) {
      if (k < g) {
        const size_t o = (size_t)(c0 + k) * rows_pad + (size_t)win * TILE_WIN + tid;
        v[k] = partial[o]; q[k] = inv[o];
      }
    }
    if (c0) __syncthreads();
#pragma unroll
    for (int k = 0; k < COMB_G; ++k) if (k < g) sp[k][tid] = v[k];
    __syncthreads();
#pragma unroll
    for (int k = 0; k < COMB_G; ++k) if (k < g) acc += sp[k][q[k]];
  }
  if (i < rows) w[i] = rh * acc;
}

static void run(const char *name, const Host &H, int cb, int ncta_total) {
  const int rows = H.rows, cols = H.cols;
  const long long nnz = (long long)H.col.size();
  const int nb = (cols + cb - 1) / cb, rows_pad = ((rows + TILE_WIN - 1) / TILE_WIN) * TILE_WIN;
  // ---- host tile build: per block, windows of 1024 rows sorted by their count in the block
  const int nwin = (rows + TILE_WIN - 1) / TILE_WIN;
  std::vector<unsigned short> perm((size_t)nb * rows_pad, 0);
  std::vector<int> sp((size_t)nb * nwin * 32 + 1, 0);
  std::vector<unsigned short> tc;
  std::vector<double> tv;
  tc.reserve(nnz + nnz / 8); tv.reserve(nnz + nnz / 8);
  int passes = 0, over = 0;
  for (int c = 0; c < nb; ++c) {
    const int lo = c * cb, hi = std::min(cols, lo + cb);
    for (int w = 0; w < nwin; ++w) {
      const int r0 = w * TILE_WIN, nr = std::min(TILE_WIN, rows - r0);
      std::vector<int> beg(nr), num(nr), ord(nr);
      for (int k = 0; k < nr; ++k) {
        const int r = r0 + k;
        const int *a = H.col.data() + H.rp[r], *b = H.col.data() + H.rp[r + 1];
        const int *q0 = std::lower_bound(a, b, lo), *q1 = std::lower_bound(a, b, hi);
        beg[k] = H.rp[r] + (int)(q0 - a); num[k] = (int)(q1 - q0); ord[k] = k;
      }
      std::stable_sort(ord.begin(), ord.end(), [&](int a, int b) { return num[a] < num[b]; });
      for (int k = 0; k < nr; ++k) perm[(size_t)c * rows_pad + r0 + ord[k]] = (unsigned short)k;
      for (int s = 0; s < 32; ++s) {
        sp[((size_t)c * nwin + w) * 32 + s] = passes;
        int maxc = 0;
        for (int l = 0; l < 32; ++l) { const int k = s * 32 + l; if (k < nr) maxc = std::max(maxc, num[ord[k]]); }
        for (int p = 0; p < maxc; ++p)
          for (int l = 0; l < 32; ++l) {
            const int k = s * 32 + l;
            if (k < nr && num[ord[k]] > p) { tc.push_back((unsigned short)(H.col[beg[ord[k]] + p] - lo)); tv.push_back(H.val[beg[ord[k]] + p]); }
            else { tc.push_back(0); tv.push_back(0.0); }
          }
        passes += maxc;
      }
    }
  }
  sp[(size_t)nb * nwin * 32] = passes;
  (void)over;
  // ---- items: (block, run of slices) per CTA, balanced by passes + a table-load charge per item
  std::vector<int> item_ptr(1, 0), item_blk, item_ws;
  {
    const int spb = nwin * 32;                           // slices per block
    const double tcost = 0.10 * passes / ncta_total;
    const double cap = ((double)passes + (ncta_total + nb) * tcost) / ncta_total;
    int cta = 0; double used = 0;
    auto add_item = [&](int c, int a, int e) {
      item_blk.push_back(c);
      for (int w = 0; w <= 32; ++w) {
        const long long target = sp[a] + (long long)(sp[e] - sp[a]) * w / 32;
        int q = (int)(std::lower_bound(sp.begin() + a, sp.begin() + e + 1, (int)target) - sp.begin());
        if (w == 32) q = e;
        item_ws.push_back(q);
      }
    };
    for (int c = 0; c < nb; ++c) {
      int pos = c * spb; const int end = (c + 1) * spb;
      while (pos < end) {
        if (used + tcost >= cap && cta < ncta_total - 1) { ++cta; item_ptr.push_back((int)item_blk.size()); used = 0; }
        const double room = cap - used - tcost;
        int e = (int)(std::upper_bound(sp.begin() + pos, sp.begin() + end + 1, sp[pos] + (int)room) - sp.begin()) - 1;
        e = std::min(std::max(e, pos + 1), end);
        if (cta == ncta_total - 1) e = end;
        add_item(c, pos, e);
        used += tcost + (sp[e] - sp[pos]);
        pos = e;
      }
    }
    while ((int)item_ptr.size() < ncta_total + 1) item_ptr.push_back((int)item_blk.size());
  }
  TileView T{};
  T.rows = rows; T.rows_pad = rows_pad; T.cols = cols; T.cb = cb; T.nblocks = nb; T.nwin = nwin; T.ncta = ncta_total;
  T.slice_ptr = up(sp); T.inv = up(perm); T.col = up(tc, 4096); T.val = up(tv, 4096);
  T.item_ptr = up(item_ptr); T.item_blk = up(item_blk); T.item_ws = up(item_ws);
  const int nitems = (int)item_blk.size();
  std::vector<double> x(cols), rho(rows);
  for (auto &t : x) t = (double)(rnd() % 2001) / 1000.0 - 1.0;
  for (auto &t : rho) t = 0.5 + (double)(rnd() % 1000) / 1000.0;
  double *dx = up(x), *drho = up(rho), *dpart, *dw;
  cudaMalloc(&dpart, (size_t)nb * rows_pad * 8); cudaMemset(dpart, 0, (size_t)nb * rows_pad * 8); cudaMalloc(&dw, (size_t)rows * 8);
  const size_t smem = tile_smem_bytes(T, 1);
  cudaFuncSetAttribute(spmv_tile_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const float tt = timeit([&] { spmv_tile_kernel<1><<<T.ncta, TILE_THREADS, smem>>>(T, dx, nullptr, dpart, nullptr); });
  const float tc2 = timeit([&] { k_combine<<<nwin, 1024>>>(dpart, T.inv, nb, rows, rows_pad, drho, dw); });
  const float tb = timeit([&] {
    spmv_tile_kernel<1><<<T.ncta, TILE_THREADS, smem>>>(T, dx, nullptr, dpart, nullptr);
    k_combine<<<nwin, 1024>>>(dpart, T.inv, nb, rows, rows_pad, drho, dw);
  });
  cudaError_t err = cudaDeviceSynchronize();
  std::vector<double> w(rows);
  cudaMemcpy(w.data(), dw, (size_t)rows * 8, cudaMemcpyDeviceToHost);
  double dmax = 0, amax = 0;
  for (int r = 0; r < rows; ++r) {
    double s = 0;
    for (int k = H.rp[r]; k < H.rp[r + 1]; ++k) s += H.val[k] * x[H.col[k]];
    s *= rho[r];
    dmax = std::max(dmax, std::fabs(s - w[r])); amax = std::max(amax, std::fabs(s));
  }
  const double mb12 = (12.0 * nnz + 4.0 * rows + 8.0 * cols + 16.0 * rows) / 1e6;   // SURVEY 8(d) bytes, as spmv_sell_bench
  const double mbt = (10.0 * 32 * passes + (double)nb * rows_pad * 10 + 8.0 * cols) / 1e6;   // bytes the tile kernel really moves
  printf("%-3s ch=%d ld=%d cb=%d blocks=%d items=%d pad=%.3f | tile %7.2f us (%5.0f GB/s real, %5.0f GB/s algorithmic)  "
         "combine %6.2f us  both %7.2f us (%5.0f GB/s algorithmic)  maxdiff %.2e (max|w| %.2e)  %s\n",
         name, TILE_CH, TILE_LD, cb, nb, nitems, 32.0 * passes / nnz, tt, mbt / tt * 1e3, mb12 / tt * 1e3, tc2, tb, mb12 / tb * 1e3,
         dmax, amax, cudaGetErrorString(err));
  fflush(stdout);
}

int main(int argc, char **argv) {
  const int n = 100000, m = 1000000;
  const int which = argc > 1 ? atoi(argv[1]) : 3;
  if (which & 1) {
    Host A; A.rows = m; A.cols = n; A.rp.resize(m + 1);
    A.col.resize((size_t)m * 10); A.val.resize((size_t)m * 10);
    for (int r = 0; r <= m; ++r) A.rp[r] = r * 10;
    for (int r = 0; r < m; ++r) {
      int *c = &A.col[(size_t)r * 10];
      for (int k = 0; k < 10; ++k) c[k] = (int)(rnd() % n);
      std::sort(c, c + 10);
      for (int k = 0; k < 10; ++k) A.val[(size_t)r * 10 + k] = (double)(rnd() % 2001) / 1000.0 - 1.0;
    }
    for (int cb : {25000, 20000, 16668, 12500}) run("A", A, cb, 148);
  }
  if (which & 2) {
    Host M; M.rows = n; M.cols = m; M.rp.resize(n + 1); M.rp[0] = 0;
    for (int r = 0; r < n; ++r) {
      const int la = 90 + (int)(rnd() % 21);
      std::vector<int> c2;
      for (int k = 0; k < la; ++k) c2.push_back((int)(rnd() % m));
      std::sort(c2.begin(), c2.end());
      for (int cc : c2) { M.col.push_back(cc); M.val.push_back((double)(rnd() % 2001) / 1000.0 - 1.0); }
      M.rp[r + 1] = (int)M.col.size();
    }
    for (int cb : {26316, 23256, 20000, 16000}) run("At", M, cb, 148);
  }
  return 0;
}
