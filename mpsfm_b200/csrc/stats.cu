// stats.cu -- problem-assembly arithmetic and order statistics (SURVEY.md section 8, row f "next-2").
//
// mpba_depth_prior_terms: the per-image numpy of Optimizer.__build_problem
// (/root/reference/mpsfm/sfm/mapper/bundle_adjustment.py:136-161), __build_shiftscale_problem (:239-241) and
// update_truncation_multiplier (:311-330) for every sampled observation of a bundle in one launch.
//
// mpba_segmented_median: np.median per segment (:241, one segment per image) and fit_robust_gaussian_mad
// (:10-15, one segment over everything) as an exact radix select.  HBM-bound: every pass streams the values
// once (8 B / element); nothing is sorted or moved.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include <mutex>
#include <string>

#include "../../include/mpba.h"

namespace {

// ------------------------------------------------------------------------------------------------------------
// depth-prior terms
// ------------------------------------------------------------------------------------------------------------
struct TermsArgs {
  int n;
  const int32_t* img_of;
  const double* row3;
  const double* xyz;
  const double* depth;
  const double* variance;
  double f, inv_f, m;
  int gross;
  double* z;
  uint8_t* mask;
  double* w;
  double* a;
  double* logratio;
  double* whitened;
};

__global__ void __launch_bounds__(256) k_depth_prior_terms(const TermsArgs A) {
  const int k = blockIdx.x * 256 + threadIdx.x;
  if (k >= A.n) return;
  const int im = __ldg(A.img_of + k);
  const double2 r01 = __ldg(reinterpret_cast<const double2*>(A.row3) + 2 * im);
  const double2 r23 = __ldg(reinterpret_cast<const double2*>(A.row3) + 2 * im + 1);
  const double x = __ldg(A.xyz + 3 * (size_t)k), y = __ldg(A.xyz + 3 * (size_t)k + 1), zz = __ldg(A.xyz + 3 * (size_t)k + 2);
  const double z = r01.x * x + r01.y * y + r23.x * zz + r23.y;
  const double d = __ldg(A.depth + k), v = __ldg(A.variance + k);
  const double sv = sqrt(v);
  const double log_d = log(d), log_z = log(z);
  if (A.z) A.z[k] = z;
  if (A.mask) {
    bool ok = d > 0.0;
    if (A.f > 0.0) {
      const double div = __ddiv_rn(d, z);
      ok = ok && div < A.f && div > A.inv_f;
    }
    if (A.gross) {
      const double wh = __ddiv_rn(fabs(fmax(log_d, 1e-6) - fmax(log_z, 1e-6)), sv);
      ok = ok && wh < 3.0;
    }
    A.mask[k] = ok ? 1 : 0;
  }
  if (A.w) A.w[k] = __dmul_rn(__dmul_rn(d, d), __ddiv_rn(1.0, fmax(v, 1e-6)));
  if (A.a) A.a[k] = __ddiv_rn(__dmul_rn(A.m, sv), d);
  if (A.logratio) A.logratio[k] = log(fmax(__ddiv_rn(z, d), 1e-6));
  if (A.whitened) A.whitened[k] = __ddiv_rn(__dadd_rn(log_d, -log_z), fmax(__ddiv_rn(sv, d), 1e-6));
}

// ------------------------------------------------------------------------------------------------------------
// segmented radix select
// ------------------------------------------------------------------------------------------------------------
constexpr int SEL_THREADS = 256;
constexpr int SEL_PER_THREAD = 16;
constexpr int SEL_CHUNK = SEL_THREADS * SEL_PER_THREAD;  // elements of one segment handled by one CTA

struct SelState {
  unsigned long long prefix[2];  // key bits fixed so far, for rank (n-1)/2 and n/2
  uint32_t rank[2];              // rank among the elements that match the prefix
  uint32_t has_nan;
  uint32_t pad;
};

struct SelWork {
  int32_t* chunk_ptr;  // [S+1] first CTA of every segment
  uint32_t* hist;      // [S][2][256]
  SelState* state;     // [S]
  double* dev;         // [n] |x - median| (MAD pass)
};

__device__ __forceinline__ unsigned long long order_key(double x) {
  if (x != x) return ~0ull;  // NaNs sort last, as in numpy
  const unsigned long long u = (unsigned long long)__double_as_longlong(x);
  return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}

__device__ __forceinline__ double key_value(unsigned long long k) {
  const unsigned long long u = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
  return __longlong_as_double((long long)u);
}

// one CTA: chunk_ptr = exclusive scan of ceil(len / SEL_CHUNK); states and histograms cleared
__global__ void __launch_bounds__(1024) k_sel_init(int S, const int32_t* __restrict__ seg_ptr, SelWork W) {
  __shared__ int s_warp[32];
  __shared__ int s_carry;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  for (int base = 0; base < S; base += 1024) {
    const int s = base + tid;
    int len = 0;
    if (s < S) len = max(seg_ptr[s + 1] - seg_ptr[s], 0);
    const int cnt = (len + SEL_CHUNK - 1) / SEL_CHUNK;
    int inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      int v = s_warp[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
      }
      s_warp[lane] = v;
    }
    __syncthreads();
    const int before = s_carry + (warp ? s_warp[warp - 1] : 0) + inc - cnt;
    if (s < S) {
      W.chunk_ptr[s] = before;
      SelState st;
      st.prefix[0] = st.prefix[1] = 0ull;
      st.rank[0] = len ? (uint32_t)((len - 1) / 2) : 0u;
      st.rank[1] = (uint32_t)(len / 2);
      st.has_nan = 0u;
      st.pad = 0u;
      W.state[s] = st;
    }
    __syncthreads();
    if (tid == 1023) s_carry = before + cnt;
    __syncthreads();
  }
  if (tid == 0) W.chunk_ptr[S] = s_carry;
  for (size_t i = tid; i < (size_t)S * 512; i += 1024) W.hist[i] = 0u;
}

// CTA -> (segment, chunk within the segment); false if the CTA has nothing to do
__device__ __forceinline__ bool sel_locate(int S, const int32_t* __restrict__ seg_ptr, const int32_t* __restrict__ chunk_ptr,
                                           int& seg, int& begin, int& end) {
  const int b = blockIdx.x;
  if (b >= chunk_ptr[S]) return false;
  int lo = 0, hi = S;  // last s with chunk_ptr[s] <= b
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (chunk_ptr[mid] <= b) lo = mid; else hi = mid;
  }
  seg = lo;  // chunk_ptr[lo] <= b < chunk_ptr[lo + 1]: segments without elements own no CTA
  begin = seg_ptr[seg] + (b - chunk_ptr[seg]) * SEL_CHUNK;
  end = min(begin + SEL_CHUNK, seg_ptr[seg + 1]);
  return true;
}

// histogram of byte `pass` (0 = most significant) over the elements that match each rank's prefix
__global__ void __launch_bounds__(SEL_THREADS) k_sel_hist(int S, const int32_t* __restrict__ seg_ptr,
                                                          const double* __restrict__ values, SelWork W, int pass) {
  __shared__ uint32_t s_hist[2][257];  // bin 256: no match
  __shared__ int s_seg, s_begin, s_end;
  const int tid = threadIdx.x;
  if (tid == 0) {
    int seg = -1, b = 0, e = 0;
    if (!sel_locate(S, seg_ptr, W.chunk_ptr, seg, b, e)) seg = -1;
    s_seg = seg; s_begin = b; s_end = e;
  }
  for (int i = tid; i < 2 * 257; i += SEL_THREADS) (&s_hist[0][0])[i] = 0u;
  __syncthreads();
  const int seg = s_seg;
  if (seg < 0) return;
  const int begin = s_begin, end = s_end;
  const SelState st = W.state[seg];
  const int shift = 56 - 8 * pass;
  const unsigned long long hi_mask = pass ? (~0ull << (shift + 8)) : 0ull;
  const bool same = st.prefix[0] == st.prefix[1];
  bool nan_seen = false;
#pragma unroll 4
  for (int j = 0; j < SEL_PER_THREAD; ++j) {
    const int i = begin + j * SEL_THREADS + tid;
    int b0 = 256, b1 = 256;
    if (i < end) {
      const double x = __ldg(values + i);
      nan_seen |= (x != x);
      const unsigned long long key = order_key(x);
      const int byte = (int)((key >> shift) & 0xffull);
      if ((key & hi_mask) == st.prefix[0]) b0 = byte;
      if (!same && (key & hi_mask) == st.prefix[1]) b1 = byte;
    }
    // one shared-memory atomic per distinct bin and warp (the top bytes of a segment are nearly constant)
    const unsigned peers0 = __match_any_sync(0xffffffffu, b0);
    if (b0 != 256 && (threadIdx.x & 31) == __ffs(peers0) - 1) atomicAdd(&s_hist[0][b0], (uint32_t)__popc(peers0));
    if (!same) {
      const unsigned peers1 = __match_any_sync(0xffffffffu, b1);
      if (b1 != 256 && (threadIdx.x & 31) == __ffs(peers1) - 1) atomicAdd(&s_hist[1][b1], (uint32_t)__popc(peers1));
    }
  }
  if (pass == 0 && nan_seen) W.state[seg].has_nan = 1u;  // benign race: every writer stores 1
  __syncthreads();
  uint32_t* gh = W.hist + (size_t)seg * 512;
  {
    const uint32_t c0 = s_hist[0][tid];
    if (c0) atomicAdd(gh + tid, c0);
    const uint32_t c1 = same ? c0 : s_hist[1][tid];
    if (c1) atomicAdd(gh + 256 + tid, c1);
  }
}

// one warp per (segment, rank): bucket that holds the rank, narrow the prefix, clear the histogram
__global__ void __launch_bounds__(256) k_sel_pick(int S, SelWork W, int pass) {
  const int gw = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (gw >= 2 * S) return;
  const int seg = gw >> 1, which = gw & 1;
  uint32_t* gh = W.hist + (size_t)seg * 512 + which * 256;
  uint32_t c[8];
  uint32_t sum = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    c[j] = gh[lane * 8 + j];
    sum += c[j];
  }
  uint32_t inc = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  const uint32_t rank = W.state[seg].rank[which];
  const unsigned holders = __ballot_sync(0xffffffffu, inc > rank);
#pragma unroll
  for (int j = 0; j < 8; ++j) gh[lane * 8 + j] = 0u;
  if (holders == 0u) return;  // empty segment
  if (lane == __ffs(holders) - 1) {
    uint32_t before = inc - sum;
    int j = 0;
    while (j < 7 && before + c[j] <= rank) before += c[j++];
    W.state[seg].prefix[which] |= (unsigned long long)(lane * 8 + j) << (56 - 8 * pass);
    W.state[seg].rank[which] = rank - before;
  }
}

// out[s] = mean of the two selected order statistics; state reset for a following MAD select
__global__ void __launch_bounds__(256) k_sel_finish(int S, const int32_t* __restrict__ seg_ptr, SelWork W,
                                                    double* __restrict__ out) {
  const int s = blockIdx.x * 256 + threadIdx.x;
  if (s >= S) return;
  const int len = seg_ptr[s + 1] - seg_ptr[s];
  SelState st = W.state[s];
  double r;
  if (len <= 0 || st.has_nan) {
    r = __longlong_as_double(0x7ff8000000000000ll);
  } else {
    const double lo = key_value(st.prefix[0]), hi = key_value(st.prefix[1]);
    r = (len & 1) ? lo : __dmul_rn(__dadd_rn(lo, hi), 0.5);
  }
  out[s] = r;
  st.prefix[0] = st.prefix[1] = 0ull;
  st.rank[0] = len > 0 ? (uint32_t)((len - 1) / 2) : 0u;
  st.rank[1] = len > 0 ? (uint32_t)(len / 2) : 0u;
  W.state[s] = st;  // has_nan is kept: the deviations of a NaN segment are NaN as well
}

__global__ void __launch_bounds__(SEL_THREADS) k_abs_dev(int S, const int32_t* __restrict__ seg_ptr,
                                                         const double* __restrict__ values,
                                                         const double* __restrict__ median, SelWork W) {
  __shared__ int s_seg, s_begin, s_end;
  if (threadIdx.x == 0) {
    int seg = -1, b = 0, e = 0;
    if (!sel_locate(S, seg_ptr, W.chunk_ptr, seg, b, e)) seg = -1;
    s_seg = seg; s_begin = b; s_end = e;
  }
  __syncthreads();
  if (s_seg < 0) return;
  const double mu = median[s_seg];
#pragma unroll 4
  for (int j = 0; j < SEL_PER_THREAD; ++j) {
    const int i = s_begin + j * SEL_THREADS + threadIdx.x;
    if (i < s_end) W.dev[i] = fabs(__dadd_rn(__ldg(values + i), -mu));
  }
}

inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

SelWork carve_sel(void* ws, int n, int S, size_t* total) {
  char* p = static_cast<char*>(ws);
  size_t off = 0;
  SelWork W;
  W.chunk_ptr = reinterpret_cast<int32_t*>(p + off); off += align256(sizeof(int32_t) * ((size_t)S + 1));
  W.hist = reinterpret_cast<uint32_t*>(p + off);     off += align256(sizeof(uint32_t) * 512 * (size_t)S);
  W.state = reinterpret_cast<SelState*>(p + off);    off += align256(sizeof(SelState) * (size_t)S);
  W.dev = reinterpret_cast<double*>(p + off);        off += align256(sizeof(double) * (size_t)n);
  if (total) *total = off;
  return W;
}

// the 8 byte passes of one select over `values`
int run_select(int n, int S, const int32_t* seg_ptr, const double* values, const SelWork& W, cudaStream_t st) {
  const int max_ctas = n / SEL_CHUNK + S + 1;  // sum of ceil(len_s / SEL_CHUNK) <= n / SEL_CHUNK + S
  for (int pass = 0; pass < 8; ++pass) {
    k_sel_hist<<<max_ctas, SEL_THREADS, 0, st>>>(S, seg_ptr, values, W, pass);
    k_sel_pick<<<(2 * S * 32 + 255) / 256, 256, 0, st>>>(S, W, pass);
  }
  return cudaGetLastError() == cudaSuccess ? MPBA_OK : MPBA_E_CUDA;
}

}  // namespace

extern "C" int mpba_depth_prior_terms(mpba_handle* h, int32_t n, const int32_t* img_of, int32_t num_images,
                                      const double* row3, const double* xyz, const double* depth,
                                      const double* variance, double scale_filter_factor, int32_t gross_outliers,
                                      double loss_multiplier, double* out_z, uint8_t* out_mask, double* out_w,
                                      double* out_a, double* out_logratio, double* out_whitened, void* cuda_stream) {
  (void)h;
  if (n < 0 || num_images < 0) return MPBA_E_INVALID;
  if (n == 0) return MPBA_OK;
  if (!img_of || !row3 || !xyz || !depth || !variance || num_images == 0) return MPBA_E_INVALID;
  TermsArgs A;
  A.n = n; A.img_of = img_of; A.row3 = row3; A.xyz = xyz; A.depth = depth; A.variance = variance;
  A.f = scale_filter_factor > 0.0 ? scale_filter_factor : 0.0;
  A.inv_f = scale_filter_factor > 0.0 ? 1.0 / scale_filter_factor : 0.0;
  A.m = loss_multiplier; A.gross = gross_outliers ? 1 : 0;
  A.z = out_z; A.mask = out_mask; A.w = out_w; A.a = out_a; A.logratio = out_logratio; A.whitened = out_whitened;
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  k_depth_prior_terms<<<(n + 255) / 256, 256, 0, st>>>(A);
  if (cudaGetLastError() != cudaSuccess) return MPBA_E_CUDA;
  if (cudaStreamSynchronize(st) != cudaSuccess) return MPBA_E_CUDA;
  return MPBA_OK;
}

extern "C" int mpba_segmented_median_workspace_bytes(int32_t n, int32_t num_segments, size_t* bytes) {
  if (n < 0 || num_segments < 0 || !bytes) return MPBA_E_INVALID;
  size_t total = 0;
  carve_sel(nullptr, n, num_segments, &total);
  *bytes = total;
  return MPBA_OK;
}

extern "C" int mpba_segmented_median(mpba_handle* h, int32_t n, const double* values, int32_t num_segments,
                                     const int32_t* seg_ptr, void* workspace, size_t workspace_bytes,
                                     double* out_median, double* out_mad, void* cuda_stream) {
  (void)h;
  if (n < 0 || num_segments < 0) return MPBA_E_INVALID;
  if (num_segments == 0) return MPBA_OK;
  if (!seg_ptr || !workspace || !out_median || (n > 0 && !values)) return MPBA_E_INVALID;
  if (reinterpret_cast<uintptr_t>(workspace) & 255) return MPBA_E_INVALID;
  size_t need = 0;
  const int S = num_segments;
  SelWork W = carve_sel(workspace, n, S, &need);
  if (workspace_bytes < need) return MPBA_E_WORKSPACE;
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  k_sel_init<<<1, 1024, 0, st>>>(S, seg_ptr, W);
  if (run_select(n, S, seg_ptr, values, W, st) != MPBA_OK) return MPBA_E_CUDA;
  k_sel_finish<<<(S + 255) / 256, 256, 0, st>>>(S, seg_ptr, W, out_median);
  if (out_mad) {
    k_abs_dev<<<n / SEL_CHUNK + S + 1, SEL_THREADS, 0, st>>>(S, seg_ptr, values, out_median, W);
    if (run_select(n, S, seg_ptr, W.dev, W, st) != MPBA_OK) return MPBA_E_CUDA;
    k_sel_finish<<<(S + 255) / 256, 256, 0, st>>>(S, seg_ptr, W, out_mad);
  }
  if (cudaGetLastError() != cudaSuccess) return MPBA_E_CUDA;
  if (cudaStreamSynchronize(st) != cudaSuccess) return MPBA_E_CUDA;
  return MPBA_OK;
}

// ------------------------------------------------------------------------------------------------------------
// mpba_segmented_median_static: the same select with its scratch in a file-scope device array -- for callers
// that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in
// the reference.
//   * g_sel_static is a module-lifetime allocation of MPBA_MEDIAN_STATIC_BYTES (~2.6 MB) on EVERY GPU that loads the
//     library, held until the process exits;
//   * in host code the NAME g_sel_static is a host-side stand-in, not a device address: the device address is taken
//     once per device, with that device current, by cudaGetSymbolAddress and cached (sel_static_base);
//   * all calls share the array and it keeps the previous call's contents: k_sel_init clears everything that must
//     start at zero (segment states, histograms), |x - median| is fully rewritten before it is read.  Calls that
//     overlap on the GPU (two streams, two host threads) RACE on it: serialise them.
// ------------------------------------------------------------------------------------------------------------
namespace {

constexpr size_t sel_align256(size_t x) { return (x + 255) & ~(size_t)255; }
constexpr size_t SEL_STATIC_BYTES =
    sel_align256(sizeof(int32_t) * ((size_t)MPBA_MEDIAN_STATIC_MAX_SEGMENTS + 1)) +
    sel_align256(sizeof(uint32_t) * 512 * (size_t)MPBA_MEDIAN_STATIC_MAX_SEGMENTS) +
    sel_align256(sizeof(SelState) * (size_t)MPBA_MEDIAN_STATIC_MAX_SEGMENTS) +
    sel_align256(sizeof(double) * (size_t)MPBA_MEDIAN_STATIC_MAX_VALUES);

static __device__ __align__(256) unsigned char g_sel_static[SEL_STATIC_BYTES];

constexpr int SEL_STATIC_MAX_DEVICES = 64;
std::mutex g_sel_static_mutex;
void* g_sel_static_addr[SEL_STATIC_MAX_DEVICES] = {nullptr};

// device address of g_sel_static on the CURRENT device (looked up once per device); nullptr + *err on failure
void* sel_static_base(cudaError_t* err) {
  int dev = -1;
  *err = cudaGetDevice(&dev);
  if (*err != cudaSuccess) return nullptr;
  if (dev < 0 || dev >= SEL_STATIC_MAX_DEVICES) {
    *err = cudaErrorInvalidDevice;
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(g_sel_static_mutex);
  if (g_sel_static_addr[dev] == nullptr) {
    void* ptr = nullptr;
    *err = cudaGetSymbolAddress(&ptr, g_sel_static);
    if (*err != cudaSuccess || ptr == nullptr) {
      if (*err == cudaSuccess) *err = cudaErrorInvalidSymbol;
      return nullptr;
    }
    g_sel_static_addr[dev] = ptr;
  }
  return g_sel_static_addr[dev];
}

}  // namespace

namespace mpba {
int handle_fail(mpba_handle* h, int code, const std::string& msg);  // ba_solver.cu: records the message (h may be NULL)
}

extern "C" int mpba_segmented_median_static(mpba_handle* h, int32_t n, const double* values, int32_t num_segments,
                                            const int32_t* seg_ptr, double* out_median, double* out_mad,
                                            void* cuda_stream) {
  if (n < 0 || num_segments < 0) return mpba::handle_fail(h, MPBA_E_INVALID, "segmented_median_static: negative size");
  if (n > MPBA_MEDIAN_STATIC_MAX_VALUES || num_segments > MPBA_MEDIAN_STATIC_MAX_SEGMENTS)
    return mpba::handle_fail(h, MPBA_E_WORKSPACE,
                             "segmented_median_static: " + std::to_string(n) + " values / " +
                                 std::to_string(num_segments) + " segments exceed the static scratch (at most " +
                                 std::to_string(MPBA_MEDIAN_STATIC_MAX_VALUES) + " values, " +
                                 std::to_string(MPBA_MEDIAN_STATIC_MAX_SEGMENTS) +
                                 " segments): use mpba_segmented_median with a caller-owned workspace");
  if (num_segments == 0) return MPBA_OK;
  if (!seg_ptr || !out_median || (n > 0 && !values))
    return mpba::handle_fail(h, MPBA_E_INVALID, "segmented_median_static: NULL array");
  cudaError_t ce = cudaSuccess;
  void* base = sel_static_base(&ce);
  if (base == nullptr)
    return mpba::handle_fail(h, MPBA_E_CUDA, std::string("segmented_median_static: cudaGetSymbolAddress: ") +
                                                 cudaGetErrorString(ce));
  const int S = num_segments;
  size_t need = 0;
  // carved for the MAXIMUM sizes, so that the layout (and the address of every part) never depends on the call
  SelWork W = carve_sel(base, MPBA_MEDIAN_STATIC_MAX_VALUES, MPBA_MEDIAN_STATIC_MAX_SEGMENTS, &need);
  if (need > SEL_STATIC_BYTES) return mpba::handle_fail(h, MPBA_E_WORKSPACE, "segmented_median_static: scratch layout");
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  k_sel_init<<<1, 1024, 0, st>>>(S, seg_ptr, W);  // clears the states and histograms of the S segments in use
  if (run_select(n, S, seg_ptr, values, W, st) != MPBA_OK) return mpba::handle_fail(h, MPBA_E_CUDA, "segmented_median_static: launch");
  k_sel_finish<<<(S + 255) / 256, 256, 0, st>>>(S, seg_ptr, W, out_median);
  if (out_mad) {
    k_abs_dev<<<n / SEL_CHUNK + S + 1, SEL_THREADS, 0, st>>>(S, seg_ptr, values, out_median, W);
    if (run_select(n, S, seg_ptr, W.dev, W, st) != MPBA_OK) return mpba::handle_fail(h, MPBA_E_CUDA, "segmented_median_static: launch");
    k_sel_finish<<<(S + 255) / 256, 256, 0, st>>>(S, seg_ptr, W, out_mad);
  }
  ce = cudaGetLastError();
  if (ce != cudaSuccess) return mpba::handle_fail(h, MPBA_E_CUDA, std::string("segmented_median_static: ") + cudaGetErrorString(ce));
  ce = cudaStreamSynchronize(st);  // same contract as mpba_segmented_median: returns after the stream completed
  if (ce != cudaSuccess) return mpba::handle_fail(h, MPBA_E_CUDA, std::string("segmented_median_static: ") + cudaGetErrorString(ce));
  return MPBA_OK;
}
