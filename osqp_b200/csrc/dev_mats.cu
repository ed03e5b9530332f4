// dev_mats.cu -- CSC (host, 64-bit) -> CSR(A) and CSR([Pf | A']) on the device.
// Format conversion is setup work: the stable radix sort is CUB (a library sort,
// like the reference's use of libc); every solve-path kernel is hand-written.
#include "dev_mats.cuh"
#include <cub/cub.cuh>
#include <algorithm>

// ------------------------------------------------------------ allocation cache
// Exact-size free lists.  A block is parked only after the device is idle (the same
// guarantee cudaFree gives), so a block handed to another engine / stream is never still in
// use.  OSQP_B200_NO_ALLOC_CACHE=1 turns the cache off; at most 32 GB are kept parked.
#include <mutex>
#include <unordered_map>
namespace {
struct DevCache {
  std::mutex mu;
  std::unordered_multimap<size_t, void *> parked;
  std::unordered_map<void *, size_t> live;
  size_t parked_bytes = 0;
  std::unordered_multimap<size_t, void *> pending;   // released, not yet known to be idle
  int enabled = -1;
};
DevCache g_cache;
}  // namespace

// Freed buffers are parked WITHOUT synchronising: they wait in `pending` until the next allocation that
// wants one of them, which synchronises the device once and moves them all to `parked` (what cudaFree
// guarantees -- nothing in flight touches the memory -- is needed at reuse, not at release: an engine
// teardown used to pay one device synchronisation per buffer).
int dev_cache_alloc(void **p, size_t bytes) {
  DevCache &C = g_cache;
  bytes = (bytes + 255) & ~(size_t)255;
  {
    std::unique_lock<std::mutex> lk(C.mu);
    if (C.enabled < 0) {
      const char *e = getenv("OSQP_B200_NO_ALLOC_CACHE");
      C.enabled = !(e && e[0] == '1');
    }
    auto it = C.parked.find(bytes);
    if (it == C.parked.end() && C.pending.find(bytes) != C.pending.end()) {
      std::unordered_multimap<size_t, void *> idle;
      idle.swap(C.pending);               // exactly what was released BEFORE the synchronisation below
      lk.unlock();
      cudaDeviceSynchronize();
      lk.lock();
      for (auto &kv : idle) C.parked.emplace(kv.first, kv.second);
      it = C.parked.find(bytes);
    }
    if (it != C.parked.end()) {
      *p = it->second;
      C.parked.erase(it);
      C.parked_bytes -= bytes;
      C.live[*p] = bytes;
      return 0;
    }
  }
  cudaError_t err = cudaMalloc(p, bytes);
  if (err != cudaSuccess) {
    // give the parked memory back and retry once
    std::unordered_multimap<size_t, void *> drop, drop2;
    {
      std::lock_guard<std::mutex> lk(C.mu);
      drop.swap(C.parked);
      drop2.swap(C.pending);
      C.parked_bytes = 0;
    }
    cudaGetLastError();
    for (auto &kv : drop) cudaFree(kv.second);
    for (auto &kv : drop2) cudaFree(kv.second);
    err = cudaMalloc(p, bytes);
    if (err != cudaSuccess) { *p = nullptr; return 1; }
  }
  std::lock_guard<std::mutex> lk(C.mu);
  C.live[*p] = bytes;
  return 0;
}

void dev_cache_free(void *p) {
  if (!p) return;
  DevCache &C = g_cache;
  size_t bytes = 0;
  {
    std::lock_guard<std::mutex> lk(C.mu);
    auto it = C.live.find(p);
    if (it != C.live.end()) { bytes = it->second; C.live.erase(it); }
    if (bytes && C.enabled == 1 && C.parked_bytes + bytes <= ((size_t)32 << 30)) {
      C.pending.emplace(bytes, p);
      C.parked_bytes += bytes;
      return;
    }
  }
  cudaFree(p);
}

// ------------------------------------------------------------------ kernels
__global__ void k_i64_to_i32(const long long *src, int *dst, long long count, int *err) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count;
       i += (long long)gridDim.x * blockDim.x) {
    const long long v = src[i];
    if (v < 0 || v > INT_MAX) *err = 1;
    dst[i] = (int)v;
  }
}
__global__ void k_iota(int *dst, int count) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    dst[i] = i;
}
// column of every CSC slot: last j with colptr[j] <= k
__global__ void k_col_of_slot(const int *colptr, int ncols, int nnz, int *out) {
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += gridDim.x * blockDim.x) {
    int lo = 0, hi = ncols;  // invariant: colptr[lo] <= k < colptr[hi]
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (colptr[mid] <= k) lo = mid; else hi = mid;
    }
    out[k] = lo;
  }
}
// ptr[r] = first position p in sorted[0..count) with sorted[p] >= r   (r = 0..rows)
__global__ void k_rowptr_from_sorted(const int *sorted, int count, int rows, int *ptr) {
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r <= rows; r += gridDim.x * blockDim.x) {
    int lo = 0, hi = count;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (sorted[mid] < r) lo = mid + 1; else hi = mid;
    }
    ptr[r] = lo;
  }
}
__global__ void k_gather_i(const int *src, const int *perm, int count, int *dst) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    dst[i] = src[perm[i]];
}
// two candidate entries of the full symmetric P per stored upper-triangular slot
__global__ void k_pf_keys(const int *prow, const int *pcol, int nnzP,
                          unsigned long long *keys, int *payload, int *ndiag) {
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < nnzP; k += gridDim.x * blockDim.x) {
    const unsigned long long i = (unsigned)prow[k], j = (unsigned)pcol[k];
    keys[2 * k] = (i << 32) | j;
    payload[2 * k] = k;
    if (i != j) {
      keys[2 * k + 1] = (j << 32) | i;
    } else {
      keys[2 * k + 1] = ~0ull;   // sorts behind every real entry
      atomicAdd(ndiag, 1);
    }
    payload[2 * k + 1] = k;
  }
}
__global__ void k_pf_unpack(const unsigned long long *keys, int count, int *row, int *col) {
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < count; p += gridDim.x * blockDim.x) {
    row[p] = (int)(keys[p] >> 32);
    col[p] = (int)(keys[p] & 0xffffffffull);
  }
}
__global__ void k_m_rowptr(const int *pf_ptr, const int *a_colptr, int n, int *rowptr, int *split) {
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j <= n; j += gridDim.x * blockDim.x) {
    rowptr[j] = pf_ptr[j] + a_colptr[j];
    if (j < n) split[j] = pf_ptr[j + 1] + a_colptr[j];
  }
}
__global__ void k_m_fill_p(const int *pf_ptr, const int *pf_row, const int *pf_col,
                           const int *pf_src, int nnzPf, const int *m_rowptr, int *col,
                           int *src, int *rowidx) {
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < nnzPf; p += gridDim.x * blockDim.x) {
    const int j = pf_row[p];
    const int dst = m_rowptr[j] + (p - pf_ptr[j]);
    col[dst] = pf_col[p];
    src[dst] = pf_src[p];
    rowidx[dst] = j;
  }
}
__global__ void k_m_fill_a(const int *a_colptr, const int *a_row, const int *a_colofslot,
                           int nnzA, int n, const int *split, int *col, int *src, int *rowidx) {
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < nnzA; k += gridDim.x * blockDim.x) {
    const int j = a_colofslot[k];
    const int dst = split[j] + (k - a_colptr[j]);
    col[dst] = n + a_row[k];
    src[dst] = k;
    rowidx[dst] = j;
  }
}
// chunk_row[k] = first row r with rowptr[r] >= k * CHUNK (rows if none); last entry = rows
__global__ void k_chunk_rows(const int *rowptr, int rows, int nchunks, int *chunk_row) {
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k <= nchunks; k += gridDim.x * blockDim.x) {
    if (k == nchunks) { chunk_row[k] = rows; continue; }
    const long long target = (long long)k * SPMV_CHUNK_NNZ;
    int lo = 0, hi = rows;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if ((long long)rowptr[mid] < target) lo = mid + 1; else hi = mid;
    }
    chunk_row[k] = lo;
  }
}
__global__ void k_chunk_desc(const int *rowptr, const int *chunk_row, int nchunks, int4 *desc) {
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < nchunks; k += gridDim.x * blockDim.x) {
    const int r0 = chunk_row[k], r1 = chunk_row[k + 1];
    desc[k] = make_int4(r0, r1, rowptr[r0], rowptr[r1]);
  }
}
__global__ void k_load_vals_a(const double *raw, const int *src, int nnz, double *val) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += gridDim.x * blockDim.x)
    val[i] = raw[src[i]];
}
__global__ void k_load_vals_m(const double *praw, const double *araw, const int *src,
                              const int *col, int n, int nnz, double *val) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += gridDim.x * blockDim.x)
    val[i] = (col[i] < n) ? praw[src[i]] : araw[src[i]];
}
__global__ void k_scatter_vals(double *raw, const double *vals, const long long *idx, int count) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    raw[idx[i]] = vals[i];
}
__global__ void k_row_absmax(const double *val, const int *rowidx, const int *col, int nnz,
                             int p_only_n, double *out) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += gridDim.x * blockDim.x) {
    if (p_only_n >= 0 && col[i] >= p_only_n) continue;
    atomic_max_nonneg(out + rowidx[i], fabs(val[i]));
  }
}
// (val * e_row) * d_col : same order as reference mat_premult_diag then mat_postmult_diag
__global__ void k_scale_a(double *val, const int *rowidx, const int *col, int nnz,
                          const double *e, const double *d) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += gridDim.x * blockDim.x)
    val[i] = (val[i] * e[rowidx[i]]) * d[col[i]];
}
__global__ void k_scale_m(double *val, const int *rowidx, const int *col, int nnz, int n,
                          const double *d, const double *e) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += gridDim.x * blockDim.x) {
    const int r = rowidx[i], c = col[i];
    if (c < n) {
      // upper entry (i<=j) is (val*d_i)*d_j; its mirror must round identically
      const int a = min(r, c), b = max(r, c);
      val[i] = (val[i] * d[a]) * d[b];
    } else {
      val[i] = (val[i] * e[c - n]) * d[r];
    }
  }
}
__global__ void k_scale_m_p(double *val, const int *col, int nnz, int n, const double *c_dev) {
  const double c = *c_dev;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += gridDim.x * blockDim.x)
    if (col[i] < n) val[i] *= c;
}

// --------------------------------------------------------------- host side
#define GRID(n) grid_for((long long)(n)), OSQP_CUDA_THREADS, 0, st

static int padded(int nnz) { return ((nnz + 3) & ~3) + 4; }


static void sell_free(SellDev &S);
static void tile_free(TileDev &T);
static void csr_free(CsrDev &C) {
  dev_free(C.rowptr); dev_free(C.col); dev_free(C.val); dev_free(C.rowidx);
  dev_free(C.src); dev_free(C.split); dev_free(C.chunk_row); dev_free(C.chunk_desc);
  sell_free(C.sell);
  tile_free(C.tile);
}

static int csr_alloc(CsrDev &C, int rows, int cols, int nnz, bool with_split, cudaStream_t st) {
  C.rows = rows; C.cols = cols; C.nnz = nnz;
  const int np = padded(nnz);
  if (dev_alloc(&C.rowptr, (size_t)rows + 8) || dev_alloc(&C.col, np) ||
      dev_alloc(&C.val, np) || dev_alloc(&C.rowidx, np) || dev_alloc(&C.src, np)) return 1;
  if (with_split && dev_alloc(&C.split, (size_t)rows + 1)) return 1;
  cudaMemsetAsync(C.col, 0, sizeof(int) * np, st);   // same stream as the fill kernels
  cudaMemsetAsync(C.val, 0, sizeof(double) * np, st);   // same stream as the fill kernels
  cudaMemsetAsync(C.rowidx, 0, sizeof(int) * np, st);   // same stream as the fill kernels
  cudaMemsetAsync(C.src, 0, sizeof(int) * np, st);   // same stream as the fill kernels
  C.nchunks = nnz / SPMV_CHUNK_NNZ + 1;
  if (dev_alloc(&C.chunk_row, (size_t)C.nchunks + 1) || dev_alloc(&C.chunk_desc, (size_t)C.nchunks)) return 1;
  C.group = spmv_direct_group(nnz, rows);
  return 0;
}

// H2D of a 64-bit index array + narrowing to int32 on the device
static int upload_idx(const c_int *host, long long count, int *dst, int *err_flag,
                      cudaStream_t st) {
  if (count <= 0) return 0;
  long long *tmp = nullptr;
  if (dev_alloc(&tmp, (size_t)count)) return 1;
  CUDA_TRY(cudaMemcpyAsync(tmp, host, sizeof(long long) * count, cudaMemcpyHostToDevice, st));
  k_i64_to_i32<<<GRID(count)>>>(tmp, dst, count, err_flag);
  CUDA_TRY(cudaStreamSynchronize(st));
  dev_free(tmp);
  return 0;
}

template <class K>
static int sort_pairs(K *keys_in, K *keys_out, int *vals_in, int *vals_out, int count,
                      int end_bit, cudaStream_t st) {
  if (count <= 0) return 0;
  size_t tmp_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys_in, keys_out, vals_in, vals_out,
                                  count, 0, end_bit, st);
  void *tmp = nullptr;
  if (dev_cache_alloc(&tmp, tmp_bytes ? tmp_bytes : 1)) return 1;
  cudaError_t e = cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys_in, keys_out, vals_in,
                                                  vals_out, count, 0, end_bit, st);
  cudaStreamSynchronize(st);
  dev_cache_free(tmp);
  CUDA_TRY(e);
  return 0;
}

static int bits_for(long long maxval) {
  int b = 1;
  while ((1LL << b) <= maxval && b < 62) ++b;
  return b;
}


// ------------------------------------------------- sliced-ELL mirror (spmv_sell.cuh)
// One warp per BLOCK of 32 rows: lanes per row G = 2^lg.  Among the values whose slices stay
// within `kmax` passes (a global figure: a warp of the persistent grid should own at least
// two slices, or the contiguous split cannot balance) take the one that pads the block
// least, charging one pass per slice for its epilogue and bookkeeping.  Measured on config 2's M (rows of
// ~107, tools/microbench): G = 4 (31 passes) 50 us, G = 8 51, G = 16 52, G = 32 58 (padding),
// G = 2 (62 passes, 1.3 slices per warp) 55.  The block then holds G slices of 32/G rows.
__global__ void k_sell_plan(const int *rowptr, int rows, int nblocks, int kmax, int *blk_lg, int *blk_ns) {
  const int lane = threadIdx.x & 31;
  for (int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < nblocks; b += (gridDim.x * blockDim.x) >> 5) {
    const int r = b * 32 + lane;
    const int len = (r < rows) ? rowptr[r + 1] - rowptr[r] : 0;
    int best_lg = 5, best = INT_MAX;
    for (int lg = 0; lg <= 5; ++lg) {
      const int rps = 32 >> lg;            // rows per slice = lanes whose max is taken together
      int mx = len;
      for (int o = 1; o < rps; o <<= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      const int k = (mx + (1 << lg) - 1) >> lg;
      int tot = ((lane & (rps - 1)) == 0) ? k : 0;      // one lane per slice contributes
      int kk = k;
      for (int o = 16; o > 0; o >>= 1) {
        tot += __shfl_xor_sync(0xffffffffu, tot, o);
        kk = max(kk, __shfl_xor_sync(0xffffffffu, kk, o));
      }
      const int cost = tot + (1 << lg);          // a slice costs about one pass of bookkeeping
      if ((kk <= kmax || lg == 5) && cost < best) { best = cost; best_lg = lg; }
    }
    if (lane == 0) { blk_lg[b] = best_lg; blk_ns[b] = 1 << best_lg; }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) blk_ns[nblocks] = 0;   // the exclusive scan leaves the total there
}
// per slice: passes = ceil(longest row / G), info = (first row << 3) | lg; largest pass count
__global__ void k_sell_slices(const int *rowptr, int rows, int nblocks, const int *blk_lg,
                              const int *blk_base, int nslices, int *len, int *info, int *maxlen) {
  const int lane = threadIdx.x & 31;
  for (int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < nblocks; b += (gridDim.x * blockDim.x) >> 5) {
    const int lg = blk_lg[b], rps = 32 >> lg;
    const int r = b * 32 + lane;
    int mx = (r < rows) ? rowptr[r + 1] - rowptr[r] : 0;
    for (int o = 1; o < rps; o <<= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((lane & (rps - 1)) == 0) {
      const int s = blk_base[b] + lane / rps;
      const int k = (mx + (1 << lg) - 1) >> lg;
      len[s] = k;
      info[s] = ((b * 32 + lane) << 3) | lg;
      if (k) atomicMax(maxlen, k);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) len[nslices] = 0;
}
// one warp per slice: (pass p, lane l) <- entry p*G + l%G of row (first row + l/G), or padding
__global__ void k_sell_fill(const int *rowptr, const int *ccol, const double *cval, int rows,
                            int nslices, const int *slice_ptr, const int *slice_info, int *col,
                            double *val, int write_col) {
  const int lane = threadIdx.x & 31;
  for (int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; s < nslices;
       s += (gridDim.x * blockDim.x) >> 5) {
    const int info = slice_info[s], lg = info & 7, G = 1 << lg;
    const int r = (info >> 3) + (lane >> lg);
    const int lo = (r < rows) ? rowptr[r] : 0, hi = (r < rows) ? rowptr[r + 1] : 0;
    const int p0 = slice_ptr[s], p1 = slice_ptr[s + 1];
    for (int p = p0; p < p1; ++p) {
      const int i = lo + (p - p0) * G + (lane & (G - 1));
      const size_t o = (size_t)p * 32 + lane;
      val[o] = (i < hi) ? cval[i] : 0.0;
      if (write_col) col[o] = (i < hi) ? ccol[i] : -1;
    }
  }
}
// warp w of the persistent grid starts at the first slice whose first pass is >= w/nwarps of all
__global__ void k_sell_warp_slice(const int *slice_ptr, int nslices, int nwarps, int *warp_slice) {
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w > nwarps) return;
  const long long total = slice_ptr[nslices];
  const long long target = total * w / nwarps;
  int lo = 0, hi = nslices;      // first s with slice_ptr[s] >= target
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (slice_ptr[mid] >= target) hi = mid; else lo = mid + 1;
  }
  warp_slice[w] = (w == nwarps) ? nslices : lo;
}

static void sell_free(SellDev &S) {
  dev_free(S.slice_ptr); dev_free(S.slice_info); dev_free(S.warp_slice); dev_free(S.col); dev_free(S.val);
  S.ok = false; S.dirty = true;
}

// values of the mirror <- current CSR values (after scaling / matrix updates)
static void sell_refresh(const CsrDev &C, cudaStream_t st) {
  SellDev &S = C.sell;
  if (!S.ok) return;
  k_sell_fill<<<grid_for((long long)S.nslices * 32), OSQP_CUDA_THREADS, 0, st>>>(
      C.rowptr, C.col, C.val, C.rows, S.nslices, S.slice_ptr, S.slice_info, S.col, S.val, 0);
  S.dirty = false;
}

static int sell_scan(const int *in, int *out, int count, cudaStream_t st) {
  size_t tmp_bytes = 0;
  void *tmp = nullptr;
  cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, in, out, count, st);
  if (dev_cache_alloc(&tmp, tmp_bytes ? tmp_bytes : 1)) return 1;
  const cudaError_t e = cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, in, out, count, st);
  const cudaError_t e2 = cudaStreamSynchronize(st);     // tmp is released below
  dev_cache_free(tmp);
  return e != cudaSuccess || e2 != cudaSuccess;
}

// Build the mirror of C (structure + current values).  Leaves sell.ok = false when the layout
// does not pay: kernel not selected, empty matrix, too much padding, or one slice so long
// that a single warp would own most of the work (the CSR kernels handle those).
static int sell_build(CsrDev &C, cudaStream_t st) {
  SellDev &S = C.sell;
  sell_free(S);
  if (spmv_kind() != 2 || C.rows <= 0 || C.nnz <= 0 || C.rows >= (1 << 28)) return 0;
  S.grid = OSQP_CUDA_SMS * SELL_MINB;
  const int nwarps = S.grid * SELL_WARPS;
  const int nblocks = (C.rows + 31) / 32;
  int *blk_lg = nullptr, *blk_ns = nullptr, *blk_base = nullptr, *len = nullptr, *maxlen = nullptr;
  int rc = 1;
  do {
    if (dev_alloc(&blk_lg, (size_t)nblocks) || dev_alloc(&blk_ns, (size_t)nblocks + 1) ||
        dev_alloc(&blk_base, (size_t)nblocks + 1) || dev_alloc(&maxlen, 1) ||
        dev_alloc(&S.warp_slice, (size_t)nwarps + 1)) break;
    cudaMemsetAsync(maxlen, 0, sizeof(int), st);
    // longest slice allowed: half the passes a warp owns on average, within [8, 48]
    const int kmax = std::min(48, std::max(8, (int)(C.nnz / 32.0 / nwarps / 2.0)));
    k_sell_plan<<<grid_for((long long)nblocks * 32), OSQP_CUDA_THREADS, 0, st>>>(C.rowptr, C.rows, nblocks,
                                                                            kmax, blk_lg, blk_ns);
    if (sell_scan(blk_ns, blk_base, nblocks + 1, st)) break;
    if (cudaMemcpy(&S.nslices, blk_base + nblocks, sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess) break;
    if (dev_alloc(&len, (size_t)S.nslices + 1) || dev_alloc(&S.slice_ptr, (size_t)S.nslices + 1) ||
        dev_alloc(&S.slice_info, (size_t)S.nslices + 1)) break;
    k_sell_slices<<<grid_for((long long)nblocks * 32), OSQP_CUDA_THREADS, 0, st>>>(
        C.rowptr, C.rows, nblocks, blk_lg, blk_base, S.nslices, len, S.slice_info, maxlen);
    if (sell_scan(len, S.slice_ptr, S.nslices + 1, st)) break;
    int h[2] = {0, 0};
    if (cudaMemcpy(&h[0], S.slice_ptr + S.nslices, sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess ||
        cudaMemcpy(&h[1], maxlen, sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess) break;
    S.total_passes = h[0];
    rc = 0;
    if (getenv("OSQP_B200_TIMING"))
      fprintf(stderr, "[sell] rows %d nnz %d: %d slices, %d passes (padding %.3f), longest slice %d passes\n", C.rows,
              C.nnz, S.nslices, S.total_passes, 32.0 * S.total_passes / C.nnz, h[1]);
    const double padded = 32.0 * S.total_passes;
    const double per_warp = (double)S.total_passes / nwarps;
    if (padded > 1.3 * C.nnz + 65536.0 || h[1] > std::max(64.0, 2.0 * per_warp) ||
        padded >= (double)INT_MAX - 512) break;        // not worth it: ok stays false
    if (dev_alloc(&S.col, (size_t)S.total_passes * 32 + 256) ||
        dev_alloc(&S.val, (size_t)S.total_passes * 32 + 256)) { rc = 1; break; }
    k_sell_warp_slice<<<grid_for(nwarps + 1), OSQP_CUDA_THREADS, 0, st>>>(S.slice_ptr, S.nslices, nwarps,
                                                                    S.warp_slice);
    k_sell_fill<<<grid_for((long long)S.nslices * 32), OSQP_CUDA_THREADS, 0, st>>>(
        C.rowptr, C.col, C.val, C.rows, S.nslices, S.slice_ptr, S.slice_info, S.col, S.val, 1);
    if (cudaGetLastError() != cudaSuccess) { rc = 1; break; }
    S.ok = true; S.dirty = false;
  } while (0);
  dev_free(blk_lg); dev_free(blk_ns); dev_free(blk_base); dev_free(len); dev_free(maxlen);
  if (!S.ok) { const int keep = rc; sell_free(S); rc = keep; }
  return rc;
}

// ------------------------------------------- column-blocked mirrors (spmv_tile.cuh)
// Built on the device from the CSR arrays.  The columns come in two ranges, each cut into its own
// blocks (TileCols): [0, n_lo) in nb_lo blocks of cb_lo (the P part of M = [P | A'], gathered from
// the n-vector; empty for A) and [n_lo, ...) in blocks of cb (gathered from the second vector).
struct TileCols { int n_lo, cb_lo, nb_lo, cb; };
__device__ __forceinline__ int tile_block_of(const TileCols &K, int c, int &local) {
  if (c < K.n_lo) { const int b = c / K.cb_lo; local = c - b * K.cb_lo; return b; }
  const int d = c - K.n_lo, b = d / K.cb;
  local = d - b * K.cb;
  return K.nb_lo + b;
}
// Step 1: one thread per row walks its entries (columns ascending): cnt[b][row] = entries of
// the row inside column block b; raises *unsorted if the columns are not ascending.
__global__ void k_tile_count(const int *rowptr, const int *col, int rows, TileCols K,
                             int rows_pad, int *cnt, int *unsorted) {
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += gridDim.x * blockDim.x) {
    const int lo = rowptr[r], hi = rowptr[r + 1];
    int prev = -1, b_cur = -1, run = 0;
    for (int e = lo; e < hi; ++e) {
      const int c = col[e];
      if (c < prev) *unsorted = 1;
      prev = c;
      int local;
      const int b = tile_block_of(K, c, local);
      if (b != b_cur) {
        if (run) cnt[(size_t)b_cur * rows_pad + r] = run;
        b_cur = b; run = 0;
      }
      ++run;
    }
    if (run) cnt[(size_t)b_cur * rows_pad + r] = run;
  }
}
// Step 2 (PARTIAL mode): one CTA per (block, window of WIN rows): stable sort of the rows by their
// count in the block (key = count << 10 | row inside the window: unique -> deterministic).
// inv[b][row] = sorted position, len[slice] = longest row of each slice of 32 sorted rows.  Rows beyond
// the matrix sort last.
template <int WIN>
__global__ void __launch_bounds__(WIN) k_tile_sort(const int *cnt, int rows, int rows_pad, int nwin,
                                                    unsigned short *inv, int *len, int *maxlen) {
  typedef cub::BlockRadixSort<unsigned, WIN, 1> Sort;
  __shared__ typename Sort::TempStorage tmp;
  const int b = blockIdx.x / nwin, win = blockIdx.x - b * nwin, tid = threadIdx.x;
  const int r = win * WIN + tid;
  unsigned key[1];
  key[0] = (r < rows) ? (((unsigned)cnt[(size_t)b * rows_pad + r] << 10) | (unsigned)tid) : (0xfffffc00u | (unsigned)tid);
  Sort(tmp).Sort(key);
  // thread tid now holds the tid-th smallest key
  const unsigned row_in_win = key[0] & 1023u;
  const bool valid = (key[0] >> 10) != 0x3fffffu;
  const int c = valid ? (int)(key[0] >> 10) : 0;
  inv[(size_t)b * rows_pad + (size_t)win * WIN + row_in_win] = (unsigned short)tid;
  // ascending order: the last valid row of a slice is its longest
  const int nvalid = min(WIN, rows - win * WIN);
  if (valid && ((tid & 31) == 31 || tid == nvalid - 1)) {
    len[(size_t)blockIdx.x * (WIN / 32) + (tid >> 5)] = c;
    if (c) atomicMax(maxlen, c);
  }
}
// Step 2 (ROWS mode): one CTA per (block, CTA row range [cta_row[g], cta_row[g + 1]), <= 1024 * ITEMS
// rows): the same sort over the whole range (key = count << 14 | row inside the range).
// inv[b][row] = sorted position inside the range (cnt / inv rows are natural: rows_pad = rows),
// perm[b][(cta_soff[g]) * 32 + sorted position] = row inside the range, len per slice as above.
template <int ITEMS>
__global__ void __launch_bounds__(1024) k_tile_sort_range(const int *cnt, int rows, int ncta, const int *cta_row,
                                                           const int *cta_soff, int nsb, unsigned short *inv,
                                                           unsigned short *perm, int *len, int *maxlen) {
  typedef cub::BlockRadixSort<unsigned, 1024, ITEMS> Sort;
  __shared__ typename Sort::TempStorage tmp;
  const int b = blockIdx.x / ncta, g = blockIdx.x - b * ncta, tid = threadIdx.x;
  const int r0 = cta_row[g], nr = cta_row[g + 1] - r0, so = cta_soff[g], ns = cta_soff[g + 1] - so;
  unsigned key[ITEMS];
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const int idx = tid * ITEMS + i;
    key[i] = (idx < nr) ? (((unsigned)cnt[(size_t)b * rows + r0 + idx] << 14) | (unsigned)idx) : (0xffffc000u | (unsigned)idx);
  }
  Sort(tmp).Sort(key);
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const int pos = tid * ITEMS + i;              // sorted position (blocked arrangement)
    const unsigned local = key[i] & 0x3fffu;
    const bool valid = (key[i] >> 14) != 0x3ffffu;
    const int c = valid ? (int)(key[i] >> 14) : 0;
    if (valid) inv[(size_t)b * rows + r0 + local] = (unsigned short)pos;
    if (pos < ns * 32) perm[((size_t)b * nsb + so) * 32 + pos] = (unsigned short)(valid ? local : pos);
    if (valid && ((pos & 31) == 31 || pos == nr - 1)) {
      len[(size_t)b * nsb + so + (pos >> 5)] = c;
      if (c) atomicMax(maxlen, c);
    }
  }
}
// Step 4: one thread per row walks its entries again and drops each into its slot: lane = sorted
// position % 32 of slice (sorted position / 32) of the row's window (PARTIAL) or CTA range (ROWS),
// pass = rank of the entry among the row's entries of block b.
__global__ void k_tile_fill(const int *rowptr, const int *col, int rows, TileCols K,
                            int inv_stride, int nwin, int win, int nsb, int ncta, const int *cta_row,
                            const int *cta_soff, const unsigned short *inv, const int *slice_ptr,
                            unsigned short *tcol, int *tsrc) {
  const int spw = win >> 5;
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += gridDim.x * blockDim.x) {
    const int lo = rowptr[r], hi = rowptr[r + 1];
    int sfirst;                                   // first slice (inside a block) of the row's window / range
    if (cta_row) {
      int a = 0, e = ncta;                        // last g with cta_row[g] <= r
      while (e - a > 1) { const int mid = (a + e) >> 1; if (cta_row[mid] <= r) a = mid; else e = mid; }
      sfirst = cta_soff[a];
    } else {
      sfirst = (r / win) * spw;
    }
    int b_cur = -1, p = 0;
    size_t slot0 = 0;
    for (int e = lo; e < hi; ++e) {
      int local;
      const int b = tile_block_of(K, col[e], local);
      if (b != b_cur) {
        b_cur = b; p = 0;
        const int k = inv[(size_t)b * inv_stride + r];
        const int s = b * nsb + sfirst + (k >> 5);
        slot0 = (size_t)slice_ptr[s] * 32 + (k & 31);
      }
      const size_t slot = slot0 + (size_t)p * 32;
      tcol[slot] = (unsigned short)local;
      tsrc[slot] = e;
      ++p;
    }
  }
}
__global__ void k_tile_vals(const int *tsrc, const double *cval, size_t slots, double *tval) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < slots; i += (size_t)gridDim.x * blockDim.x) {
    const int e = tsrc[i];
    tval[i] = e >= 0 ? cval[e] : 0.0;
  }
}

static void tile_free(TileDev &T) {
  dev_free(T.slice_ptr); dev_free(T.item_ptr); dev_free(T.item_blk); dev_free(T.item_ws); dev_free(T.cta_row);
  dev_free(T.cta_soff); dev_free(T.wb_ptr); dev_free(T.wb_blk); dev_free(T.ne); dev_free(T.src); dev_free(T.perm); dev_free(T.col); dev_free(T.val); dev_free(T.partial);

  T.ok = false; T.dirty = true;
}
static void tile_refresh(const CsrDev &C, cudaStream_t st) {
  TileDev &T = C.tile;
  if (!T.ok) return;
  const size_t slots = (size_t)T.total_passes * 32;
  k_tile_vals<<<grid_for((long long)slots), OSQP_CUDA_THREADS, 0, st>>>(T.src, C.val, slots, T.val);
  T.dirty = false;
}
static int tile_env(const char *name, int dflt) {
  const char *e = getenv(name);
  return (e && *e) ? atoi(e) : dflt;
}
// read at every build (not cached): tests switch it inside one process
static bool tile_enabled() { return tile_env("OSQP_B200_TILE", 1) != 0 && spmv_kind() == 2; }

// 32 warp ranges of entries [ka, ke) of the non-empty-slice list, balanced by passes (np = first
// pass of each listed slice, one more entry at the end); appends 33 entries to `out`
static void tile_warp_split(const std::vector<int> &np, int ka, int ke, std::vector<int> &out) {
  for (int w = 0; w <= 32; ++w) {
    const long long target = np[ka] + (long long)(np[ke] - np[ka]) * w / 32;
    int q = (int)(std::lower_bound(np.begin() + ka, np.begin() + ke + 1, (int)target) - np.begin());
    if (w == 32) q = ke;
    out.push_back(q);
  }
}

// Build the column-blocked mirror of C: columns [0, n_lo) (P of M; n_lo = 0 for A) and
// [n_lo, n_lo + cols_hi), each range in its own blocks.  Leaves
// tile.ok = false when it does not pay: small or short-row matrices, too few entries per (row,
// block), unsorted columns, too much padding, or (ROWS mode) too many rows for the shared-memory
// accumulators of 148 CTAs.
static int tile_build(CsrDev &C, int n_lo, int cols_hi, long long nnzT, bool rows_mode, cudaStream_t st) {
  TileDev &T = C.tile;
  tile_free(T);
  if (!tile_enabled() || C.rows <= 0 || cols_hi <= 0 || nnzT < (1 << 20) || (double)cols_hi * 8 < 262144.0) return 0;
  const int ncta = OSQP_CUDA_SMS;
  const int rows = C.rows;
  // shared memory: the table, and in ROWS mode 10 bytes per owned row (accumulator + perm).  The
  // streams want ~40 KB of L1 for their lines in flight, so the default budgets stay below 228 KB.
  const int budget = 1024 * tile_env(rows_mode ? "OSQP_B200_TILE_ROWS_KB" : "OSQP_B200_TILE_PART_KB", rows_mode ? 188 : 195);
  std::vector<int> cta_row, cta_soff;
  int win = TILE_WIN_PART, nwin = (rows + win - 1) / win, nsb = nwin * (win / 32), racc = 0;
  long long inv_stride = (long long)nwin * win;       // rows of one block in cnt / inv
  if (rows_mode) {
    // ---- CTA g owns rows [cta_row[g], cta_row[g + 1]), balanced by nonzeros, at most TILE_ROWS_MAX each
    std::vector<int> rp((size_t)rows + 1);
    if (cudaMemcpyAsync(rp.data(), C.rowptr, sizeof(int) * ((size_t)rows + 1), cudaMemcpyDeviceToHost, st) != cudaSuccess ||
        cudaStreamSynchronize(st) != cudaSuccess) return 1;
    // greedy cut under a load cap T (nonzeros) and the row cap; the smallest T that needs <= ncta ranges
    // (matrices that mix long and very short rows -- Lasso: 1e6 rows of 51 entries, 2e5 of 2 -- end up with
    // nonzero-balanced ranges over the long rows and row-capped ranges over the short ones)
    if ((long long)ncta * TILE_ROWS_MAX < rows) return 0;        // too many rows for this mode
    auto cut = [&](long long T, std::vector<int> *out) {
      int r = 0, g = 0;
      while (r < rows && g <= ncta) {
        const int hi = std::min(rows, r + TILE_ROWS_MAX);
        const long long lim = (long long)rp[r] + T;
        int e = (int)(std::upper_bound(rp.begin() + r, rp.begin() + hi + 1, lim > INT_MAX ? INT_MAX : (int)lim) - rp.begin()) - 1;
        if (e <= r) e = r + 1;                                    // one row heavier than T
        if (out && g < ncta) (*out)[g + 1] = e;
        r = e; ++g;
      }
      return r >= rows ? g : ncta + 1;
    };
    long long lo_t = std::max<long long>(1, ((long long)rp[rows] + ncta - 1) / ncta), hi_t = std::max<long long>(rp[rows], 1);
    if (cut(hi_t, nullptr) > ncta) return 0;
    while (lo_t < hi_t) {
      const long long mid = (lo_t + hi_t) / 2;
      if (cut(mid, nullptr) <= ncta) hi_t = mid; else lo_t = mid + 1;
    }
    cta_row.assign(ncta + 1, rows);
    cta_row[0] = 0;
    cut(lo_t, &cta_row);
    cta_soff.assign(ncta + 1, 0);
    int smax = 0;
    for (int g = 0; g < ncta; ++g) {
      const int ns = (cta_row[g + 1] - cta_row[g] + 31) / 32;
      cta_soff[g + 1] = cta_soff[g] + ns;
      smax = std::max(smax, ns);
    }
    nsb = cta_soff[ncta]; racc = smax * 32; win = 32; nwin = 0;
    inv_stride = rows;
  }
  const long long table_bytes = (long long)budget - 10LL * racc;
  if (table_bytes < 64 * 1024) return 0;
  const int cb_fit = (int)(table_bytes / 8) & ~1;
  const int nb_hi = (cols_hi + cb_fit - 1) / cb_fit;
  int cb = (cols_hi + nb_hi - 1) / nb_hi;
  cb += cb & 1;                      // bulk copies want 16-byte multiples
  const int nb_lo = n_lo > 0 ? (n_lo + cb_fit - 1) / cb_fit : 0;
  int cb_lo = nb_lo ? (n_lo + nb_lo - 1) / nb_lo : 2;
  cb_lo += cb_lo & 1;
  const int nb = nb_lo + nb_hi;
  const TileCols K{n_lo, cb_lo, nb_lo, cb};
  const double per_tile_row = (double)nnzT / ((double)rows * nb);
  // worth it: enough entries per (row, block) -- in PARTIAL mode the partial sums must stay well
  // below the matrix stream
  // (ROWS mode is judged after the build instead: padding, and the table loads of the non-empty
  // (CTA, block) phases against the matrix stream)
  // PARTIAL mode with few entries per (row, block) on average only pays on structured matrices (most
  // (window, block) pairs empty: judged after the build), and only while the dense partial array is affordable
  if ((!rows_mode && per_tile_row < 1.5 && (double)nb * (double)inv_stride * 8.0 > 1.5e9) || nb > 4096 || (long long)nb * nsb >= (1LL << 27) ||
      inv_stride * nb * 8 > (4LL << 30)) return 0;
  T.rows_mode = rows_mode; T.rows = rows; T.rows_pad = (int)inv_stride; T.cols = cols_hi; T.n_lo = n_lo;
  T.cb = cb; T.cb_lo = cb_lo; T.nb_lo = nb_lo; T.nblocks = nb; T.nwin = nwin; T.win = win; T.ncta = ncta; T.racc = racc; T.nsb = nsb;
  const size_t nsl = (size_t)nb * nsb;
  int *cnt = nullptr, *len = nullptr, *flags = nullptr;
  unsigned short *inv = nullptr, *perm = nullptr;
  int rc = 1;
  auto up = [&](int **d, const std::vector<int> &h) {
    if (dev_alloc(d, h.size() + 1)) return 1;
    if (h.empty()) return 0;
    return (int)(cudaMemcpyAsync(*d, h.data(), sizeof(int) * h.size(), cudaMemcpyHostToDevice, st) != cudaSuccess);
  };
  do {
    if (dev_alloc(&cnt, (size_t)nb * inv_stride) || dev_alloc(&len, nsl + 1) || dev_alloc(&flags, 2) ||
        dev_alloc(&inv, (size_t)nb * inv_stride) || dev_alloc(&T.slice_ptr, nsl + 1)) break;
    if (rows_mode && (dev_alloc(&perm, nsl * 32) || up(&T.cta_row, cta_row) || up(&T.cta_soff, cta_soff))) break;
    cudaMemsetAsync(cnt, 0, sizeof(int) * (size_t)nb * inv_stride, st);
    cudaMemsetAsync(len, 0, sizeof(int) * (nsl + 1), st);
    cudaMemsetAsync(flags, 0, sizeof(int) * 2, st);
    cudaMemsetAsync(inv, 0, sizeof(unsigned short) * (size_t)nb * inv_stride, st);
    k_tile_count<<<grid_for(rows), OSQP_CUDA_THREADS, 0, st>>>(C.rowptr, C.col, rows, K, (int)inv_stride, cnt, flags);
    if (rows_mode)
      k_tile_sort_range<TILE_ROWS_ITEMS><<<nb * ncta, 1024, 0, st>>>(cnt, rows, ncta, T.cta_row, T.cta_soff, nsb, inv,
                                                                     perm, len, flags + 1);
    else
      k_tile_sort<TILE_WIN_PART><<<nb * nwin, TILE_WIN_PART, 0, st>>>(cnt, rows, (int)inv_stride, nwin, inv, len, flags + 1);
    if (sell_scan(len, T.slice_ptr, (int)nsl + 1, st)) break;
    int h[2] = {0, 0};
    std::vector<int> sp(nsl + 1);
    if (cudaMemcpy(h, flags, sizeof(h), cudaMemcpyDeviceToHost) != cudaSuccess ||
        cudaMemcpy(sp.data(), T.slice_ptr, sizeof(int) * (nsl + 1), cudaMemcpyDeviceToHost) != cudaSuccess) break;
    T.total_passes = sp[nsl];
    rc = 0;
    if (getenv("OSQP_B200_TIMING"))
      fprintf(stderr, "[tile] %s rows %d cols %d nnz %lld: %d blocks of %d, %d slices per block, %d passes (padding %.3f), "
              "longest slice %d, unsorted %d, smem %d KB\n", rows_mode ? "ROWS" : "PARTIAL", rows, n_lo + cols_hi, nnzT, nb,
              cb, nsb, T.total_passes, 32.0 * T.total_passes / nnzT, h[1], h[0],
              (int)((std::max(cb, nb_lo ? cb_lo : 0) * 8LL + racc * 10LL) / 1024));
    if (h[0] || 32.0 * T.total_passes > 1.25 * nnzT + 65536.0 || h[1] > 4096) break;   // not worth it: ok stays false
    const size_t slots = (size_t)T.total_passes * 32;
    if (dev_alloc(&T.col, slots + 4096) || dev_alloc(&T.val, slots + 4096) || dev_alloc(&T.src, slots + 4096)) { rc = 1; break; }
    cudaMemsetAsync(T.col, 0, sizeof(unsigned short) * (slots + 4096), st);
    cudaMemsetAsync(T.val, 0, sizeof(double) * (slots + 4096), st);
    cudaMemsetAsync(T.src, 0xff, sizeof(int) * (slots + 4096), st);
    k_tile_fill<<<grid_for(rows), OSQP_CUDA_THREADS, 0, st>>>(C.rowptr, C.col, rows, K, (int)inv_stride,
                                                              nwin, win, nsb, ncta, T.cta_row, T.cta_soff, inv,
                                                              T.slice_ptr, T.col, T.src);
    // the slices with entries, in order: what the kernels walk (empty slices cost them nothing)
    std::vector<int> ne_slice, ne_ptr;
    ne_slice.reserve(nsl / 2 + 1); ne_ptr.reserve(nsl / 2 + 2);
    for (size_t q = 0; q < nsl; ++q)
      if (sp[q + 1] > sp[q]) { ne_slice.push_back((int)q); ne_ptr.push_back(sp[q]); }
    ne_ptr.push_back(sp[nsl]);
    auto kof = [&](size_t slice) {       // first listed slice >= slice
      return (int)(std::lower_bound(ne_slice.begin(), ne_slice.end(), (int)slice) - ne_slice.begin());
    };
    {
      std::vector<int2> ne(ne_ptr.size());
      for (size_t q = 0; q < ne_slice.size(); ++q) ne[q] = make_int2(ne_ptr[q], ne_slice[q]);
      ne.back() = make_int2(ne_ptr.back(), -1);
      if (dev_alloc(&T.ne, ne.size() + 2) ||
          cudaMemcpy(T.ne, ne.data(), sizeof(int2) * ne.size(), cudaMemcpyHostToDevice) != cudaSuccess) { rc = 1; break; }
    }
    std::vector<int> item_ptr, item_blk, item_ws;
    if (rows_mode) {
      // phases = (CTA, block) pairs with entries: each costs a table load and two CTA barriers
      long long phases = 0;
      int worst = 0;
      for (int g = 0; g < ncta; ++g) {
        int mine = 0;
        for (int c = 0; c < nb; ++c) mine += sp[c * nsb + cta_soff[g + 1]] > sp[c * nsb + cta_soff[g]];
        phases += mine; worst = std::max(worst, mine);
      }
      const double table_traffic = (double)phases * 8.0 * std::max(cb, nb_lo ? cb_lo : 0);
      if (getenv("OSQP_B200_TIMING"))
        fprintf(stderr, "[tile] ROWS phases: %lld over %d CTAs (worst %d), table loads %.0f MB against a %.0f MB stream\n",
                phases, ncta, worst, table_traffic / 1e6, 10.0 * nnzT / 1e6);
      // Measured (profiles/r2_rows_budget_sweep.log): on config 2's A (10 entries per row over 7 blocks: table loads
      // 118 MB against a 100 MB stream) the ROWS kernel takes 54-56 us and the sliced-ELL kernel 51 us; on Lasso's A
      // (130 MB against 514 MB) 151 us against 301 us.  The phases (table load + two CTA barriers each) must be
      // small change next to the stream: default bar 60 % (OSQP_B200_TILE_ROWS_MAXPCT).
      const double max_ratio = 0.01 * tile_env("OSQP_B200_TILE_ROWS_MAXPCT", 60);
      if (worst > 24 || table_traffic > max_ratio * 10.0 * (double)nnzT) { rc = 0; break; }     // not worth it
      for (int g = 0; g < ncta; ++g)
        for (int c = 0; c < nb; ++c)
          tile_warp_split(ne_ptr, kof((size_t)c * nsb + cta_soff[g]), kof((size_t)c * nsb + cta_soff[g + 1]), item_ws);
    } else {
      // ---- per window, the blocks in which it has entries: the others are neither written nor read
      std::vector<int> wb_ptr((size_t)nwin + 1, 0), wb_blk;
      const int spw = win / 32;
      for (int w = 0; w < nwin; ++w) {
        for (int c = 0; c < nb; ++c)
          if (sp[(size_t)c * nsb + (size_t)(w + 1) * spw] > sp[(size_t)c * nsb + (size_t)w * spw]) wb_blk.push_back(c);
        wb_ptr[w + 1] = (int)wb_blk.size();
      }
      const double partial_traffic = 16.0 * win * (double)wb_blk.size();      // written once, read once
      if (getenv("OSQP_B200_TIMING"))
        fprintf(stderr, "[tile] PARTIAL (window, block) pairs with entries: %zu of %lld, partial sums %.0f MB against a %.0f MB stream\n",
                wb_blk.size(), (long long)nwin * nb, partial_traffic / 1e6, 10.0 * nnzT / 1e6);
      if (partial_traffic > 0.75 * 10.0 * (double)nnzT) { rc = 0; break; }    // not worth it
      if (up(&T.wb_ptr, wb_ptr) || up(&T.wb_blk, wb_blk)) { rc = 1; break; }
      // ---- items: (block, run of slices) per CTA, balanced by passes + a table-load charge per item
      if (dev_alloc(&T.partial, (size_t)nb * inv_stride)) { rc = 1; break; }
      cudaMemsetAsync(T.partial, 0, sizeof(double) * (size_t)nb * inv_stride, st);   // empty slices are never written
      item_ptr.push_back(0);
      const double total = (double)T.total_passes;
      const double tcost = 0.10 * total / ncta;
      const double cap = (total + (double)(ncta + nb) * tcost) / ncta;
      int cta = 0; double used = 0;
      for (int c = 0; c < nb; ++c) {
        int pos = c * nsb; const int end = (c + 1) * nsb;
        if (sp[end] == sp[pos]) continue;              // block without entries
        while (pos < end) {
          if (used + tcost >= cap && cta < ncta - 1) { ++cta; item_ptr.push_back((int)item_blk.size()); used = 0; }
          const double room = cap - used - tcost;
          int e = (int)(std::upper_bound(sp.begin() + pos, sp.begin() + end + 1, sp[pos] + (int)room) - sp.begin()) - 1;
          e = std::min(std::max(e, pos + 1), end);
          if (cta == ncta - 1 || sp[e] == sp[end]) e = end;
          item_blk.push_back(c);
          tile_warp_split(ne_ptr, kof((size_t)pos), kof((size_t)e), item_ws);
          used += tcost + (sp[e] - sp[pos]);
          pos = e;
        }
      }
      while ((int)item_ptr.size() < ncta + 1) item_ptr.push_back((int)item_blk.size());
    }
    T.nitems = (int)item_blk.size();
    if (up(&T.item_ptr, item_ptr) || up(&T.item_blk, item_blk) || up(&T.item_ws, item_ws) ||
        cudaStreamSynchronize(st) != cudaSuccess || cudaGetLastError() != cudaSuccess) { rc = 1; break; }
    T.perm = rows_mode ? perm : inv;
    (rows_mode ? perm : inv) = nullptr;              // kept by the mirror
    T.ok = true;
    tile_refresh(C, st);
  } while (0);
  dev_free(cnt); dev_free(len); dev_free(flags); dev_free(inv); dev_free(perm);
  if (!T.ok) { const int keep = rc; tile_free(T); rc = keep; }
  return rc;
}

void mats_sell_sync(DevMats &W, cudaStream_t st) {
  if (W.A.sell.ok && W.A.sell.dirty) sell_refresh(W.A, st);
  if (W.M.sell.ok && W.M.sell.dirty) sell_refresh(W.M, st);
  if (W.A.tile.ok && W.A.tile.dirty) tile_refresh(W.A, st);
  if (W.M.tile.ok && W.M.tile.dirty) tile_refresh(W.M, st);
}

void mats_free(DevMats &W) {
  csr_free(W.A);
  csr_free(W.M);
  dev_free(W.Araw);
  dev_free(W.Praw);
}

// OSQP_B200_TIMING=1: phase times of the device build (each mark synchronises the stream)
#include <chrono>
struct BuildTimer {
  bool on; cudaStream_t st; std::chrono::steady_clock::time_point t0;
  BuildTimer(cudaStream_t s) : on(getenv("OSQP_B200_TIMING") != nullptr), st(s), t0(std::chrono::steady_clock::now()) {}
  void mark(const char *what) {
    if (!on) return;
    cudaStreamSynchronize(st);
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    fprintf(stderr, "[timing]   mats_build %-28s %8.3f ms\n", what, ms);
  }
};

int mats_build(DevMats &W, const csc *P, const csc *A, cudaStream_t st) {
  BuildTimer tm(st);
  const long long n = P->n, m = A->m;
  const long long nnzP = P->p[n], nnzA = (A->p ? A->p[A->n] : 0);
  if (n + m >= INT_MAX / 2 || 2 * nnzP + nnzA >= INT_MAX - 16) {
    fprintf(stderr, "osqp_cuda_pcg: problem exceeds the int32 device index range\n");
    return 1;
  }
  W.n = (int)n; W.m = (int)m; W.nnzA = (int)nnzA; W.nnzP = (int)nnzP;

  int rc = 1;
  int *err = nullptr, *ndiag = nullptr;
  int *Ap = nullptr, *Ai = nullptr, *Acol = nullptr, *Pp = nullptr, *Pi = nullptr, *Pcol = nullptr;
  int *iota = nullptr, *rows_sorted = nullptr, *perm = nullptr;
  unsigned long long *k_in = nullptr, *k_out = nullptr;
  int *pay_in = nullptr, *pay_out = nullptr, *pf_row = nullptr, *pf_col = nullptr, *pf_ptr = nullptr;
  int h_err = 0, h_ndiag = 0;

  do {
    if (dev_alloc(&err, 1) || dev_alloc(&ndiag, 1)) break;
    cudaMemsetAsync(err, 0, sizeof(int), st);
    cudaMemsetAsync(ndiag, 0, sizeof(int), st);
    if (dev_alloc(&Ap, n + 1) || dev_alloc(&Ai, nnzA) || dev_alloc(&Acol, nnzA) ||
        dev_alloc(&Pp, n + 1) || dev_alloc(&Pi, nnzP) || dev_alloc(&Pcol, nnzP)) break;
    if (dev_alloc(&W.Araw, nnzA) || dev_alloc(&W.Praw, nnzP)) break;
    if (upload_idx(A->p, n + 1, Ap, err, st) || upload_idx(A->i, nnzA, Ai, err, st) ||
        upload_idx(P->p, n + 1, Pp, err, st) || upload_idx(P->i, nnzP, Pi, err, st)) break;
    if (nnzA && cudaMemcpyAsync(W.Araw, A->x, sizeof(double) * nnzA, cudaMemcpyHostToDevice, st)) break;
    if (nnzP && cudaMemcpyAsync(W.Praw, P->x, sizeof(double) * nnzP, cudaMemcpyHostToDevice, st)) break;
    tm.mark("H2D of the CSC arrays");
    if (nnzA) k_col_of_slot<<<GRID(nnzA)>>>(Ap, (int)n, (int)nnzA, Acol);
    if (nnzP) k_col_of_slot<<<GRID(nnzP)>>>(Pp, (int)n, (int)nnzP, Pcol);

    // ---- A as CSR: stable sort of the CSC slots by row keeps columns ascending
    if (csr_alloc(W.A, (int)m, (int)n, (int)nnzA, false, st)) break;
    if (dev_alloc(&iota, nnzA) || dev_alloc(&rows_sorted, nnzA) || dev_alloc(&perm, nnzA)) break;
    if (nnzA) {
      k_iota<<<GRID(nnzA)>>>(iota, (int)nnzA);
      if (sort_pairs(Ai, rows_sorted, iota, perm, (int)nnzA, bits_for(m), st)) break;
      cudaMemcpyAsync(W.A.rowidx, rows_sorted, sizeof(int) * nnzA, cudaMemcpyDeviceToDevice, st);
      cudaMemcpyAsync(W.A.src, perm, sizeof(int) * nnzA, cudaMemcpyDeviceToDevice, st);
      k_gather_i<<<GRID(nnzA)>>>(Acol, perm, (int)nnzA, W.A.col);
    }
    k_rowptr_from_sorted<<<GRID(m + 1)>>>(rows_sorted, (int)nnzA, (int)m, W.A.rowptr);
    k_chunk_rows<<<GRID(W.A.nchunks + 1)>>>(W.A.rowptr, (int)m, W.A.nchunks, W.A.chunk_row);
    k_chunk_desc<<<GRID(W.A.nchunks)>>>(W.A.rowptr, W.A.chunk_row, W.A.nchunks, W.A.chunk_desc);

    tm.mark("CSR(A)");
    // ---- full symmetric P as sorted (row, col) pairs
    const int ncand = (int)(2 * nnzP);
    if (dev_alloc(&k_in, ncand) || dev_alloc(&k_out, ncand) || dev_alloc(&pay_in, ncand) ||
        dev_alloc(&pay_out, ncand) || dev_alloc(&pf_row, ncand) || dev_alloc(&pf_col, ncand) ||
        dev_alloc(&pf_ptr, n + 1)) break;
    if (nnzP) {
      k_pf_keys<<<GRID(nnzP)>>>(Pi, Pcol, (int)nnzP, k_in, pay_in, ndiag);
      if (sort_pairs(k_in, k_out, pay_in, pay_out, ncand, 64, st)) break;
    }
    if (cudaMemcpyAsync(&h_ndiag, ndiag, sizeof(int), cudaMemcpyDeviceToHost, st)) break;
    if (cudaMemcpyAsync(&h_err, err, sizeof(int), cudaMemcpyDeviceToHost, st)) break;
    if (cudaStreamSynchronize(st) != cudaSuccess) break;
    if (h_err) { fprintf(stderr, "osqp_cuda_pcg: index out of int32 range\n"); break; }
    W.nnzPf = ncand - h_ndiag;
    if (W.nnzPf) k_pf_unpack<<<GRID(W.nnzPf)>>>(k_out, W.nnzPf, pf_row, pf_col);
    k_rowptr_from_sorted<<<GRID(n + 1)>>>(pf_row, W.nnzPf, (int)n, pf_ptr);

    tm.mark("full symmetric P");
    // ---- M = [Pf | A'] : n rows, columns >= n address the m-vector
    const int nnzM = W.nnzPf + (int)nnzA;
    if (csr_alloc(W.M, (int)n, (int)(n + m), nnzM, true, st)) break;
    W.M.split_col = (int)n;
    k_m_rowptr<<<GRID(n + 1)>>>(pf_ptr, Ap, (int)n, W.M.rowptr, W.M.split);
    if (W.nnzPf)
      k_m_fill_p<<<GRID(W.nnzPf)>>>(pf_ptr, pf_row, pf_col, pay_out, W.nnzPf, W.M.rowptr,
                                    W.M.col, W.M.src, W.M.rowidx);
    if (nnzA)
      k_m_fill_a<<<GRID(nnzA)>>>(Ap, Ai, Acol, (int)nnzA, (int)n, W.M.split, W.M.col, W.M.src,
                                 W.M.rowidx);
    k_chunk_rows<<<GRID(W.M.nchunks + 1)>>>(W.M.rowptr, (int)n, W.M.nchunks, W.M.chunk_row);
    k_chunk_desc<<<GRID(W.M.nchunks)>>>(W.M.rowptr, W.M.chunk_row, W.M.nchunks, W.M.chunk_desc);
    if (mats_load_raw(W, st)) break;
    tm.mark("M = [Pf | A'] + values");
    if (sell_build(W.A, st) || sell_build(W.M, st)) break;
    tm.mark("sliced-ELL mirrors");
    // column-blocked mirrors (spmv_tile.cuh): A in ROWS mode, the A' part of M in PARTIAL mode (the P
    // part of M must be short rows: the tile epilogue walks it with one thread per row); each is
    // built when it pays on its own, the sliced-ELL mirrors stay for every other launch
    if (W.A.sell.ok && W.M.sell.ok && nnzA > 0) {
      const char *am = getenv("OSQP_B200_TILE_A");      // "rows" | "partial" | "off"
      const bool a_rows = !(am && !strcmp(am, "partial")), a_off = am && !strcmp(am, "off");
      if (!a_off && tile_build(W.A, 0, (int)n, nnzA, a_rows, st)) break;
      if (tile_build(W.M, (int)n, (int)m, nnzM, false, st)) break;
    }
    tm.mark("column-blocked mirrors");
    if (cudaStreamSynchronize(st) != cudaSuccess) break;
    if (cudaGetLastError() != cudaSuccess) break;
    rc = 0;
  } while (0);

  if (rc) {
    cudaError_t e = cudaGetLastError();
    fprintf(stderr, "osqp_cuda_pcg: building device matrices failed (%s)\n", cudaGetErrorString(e));
  }
  dev_free(err); dev_free(ndiag); dev_free(Ap); dev_free(Ai); dev_free(Acol); dev_free(Pp);
  dev_free(Pi); dev_free(Pcol); dev_free(iota); dev_free(rows_sorted); dev_free(perm);
  dev_free(k_in); dev_free(k_out); dev_free(pay_in); dev_free(pay_out); dev_free(pf_row);
  dev_free(pf_col); dev_free(pf_ptr);
  if (rc) mats_free(W);
  return rc;
}

int mats_load_raw(DevMats &W, cudaStream_t st) {
  W.A.sell.dirty = W.M.sell.dirty = true;
  W.A.tile.dirty = W.M.tile.dirty = true;
  if (W.A.nnz) k_load_vals_a<<<GRID(W.A.nnz)>>>(W.Araw, W.A.src, W.A.nnz, W.A.val);
  if (W.M.nnz)
    k_load_vals_m<<<GRID(W.M.nnz)>>>(W.Praw, W.Araw, W.M.src, W.M.col, W.n, W.M.nnz, W.M.val);
  return cudaGetLastError() != cudaSuccess;
}

int mats_update_raw(DevMats &W, bool isP, const c_float *vals, const c_int *idx, c_int count,
                    cudaStream_t st) {
  double *raw = isP ? W.Praw : W.Araw;
  const int nnz = isP ? W.nnzP : W.nnzA;
  if (!idx) {
    if (nnz) CUDA_TRY(cudaMemcpyAsync(raw, vals, sizeof(double) * nnz, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return 0;
  }
  if (count <= 0) return 0;
  for (c_int i = 0; i < count; ++i)
    if (idx[i] < 0 || idx[i] >= nnz) return 1;
  double *dv = nullptr; long long *di = nullptr;
  if (dev_alloc(&dv, (size_t)count) || dev_alloc(&di, (size_t)count)) { dev_free(dv); return 1; }
  cudaMemcpyAsync(dv, vals, sizeof(double) * count, cudaMemcpyHostToDevice, st);
  cudaMemcpyAsync(di, idx, sizeof(long long) * count, cudaMemcpyHostToDevice, st);
  k_scatter_vals<<<GRID(count)>>>(raw, dv, di, (int)count);
  cudaError_t e = cudaStreamSynchronize(st);
  dev_free(dv); dev_free(di);
  CUDA_TRY(e);
  return 0;
}

void mats_row_absmax(const CsrDev &C, double *out, bool p_part_only, int n, cudaStream_t st) {
  if (C.nnz)
    k_row_absmax<<<GRID(C.nnz)>>>(C.val, C.rowidx, C.col, C.nnz, p_part_only ? n : -1, out);
}
void mats_scale(DevMats &W, const double *d, const double *e, cudaStream_t st) {
  W.A.sell.dirty = W.M.sell.dirty = true;
  W.A.tile.dirty = W.M.tile.dirty = true;
  if (W.A.nnz) k_scale_a<<<GRID(W.A.nnz)>>>(W.A.val, W.A.rowidx, W.A.col, W.A.nnz, e, d);
  if (W.M.nnz) k_scale_m<<<GRID(W.M.nnz)>>>(W.M.val, W.M.rowidx, W.M.col, W.M.nnz, W.n, d, e);
}
void mats_scale_P(DevMats &W, const double *c_dev, cudaStream_t st) {
  W.M.sell.dirty = true;      // (the tiled A' part does not change, but one flag keeps it simple)
  W.M.tile.dirty = true;
  if (W.M.nnz) k_scale_m_p<<<GRID(W.M.nnz)>>>(W.M.val, W.M.col, W.M.nnz, W.n, c_dev);
}
