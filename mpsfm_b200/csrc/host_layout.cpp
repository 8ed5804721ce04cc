// host_layout.cpp -- HOST side of the problem boundary: the index arrays of mpba_problem from the caller's
// natural observation list.
//
// The reference hands Ceres one residual block per observation (COLMAP's AddImageToProblem, called from
// /root/reference/mpsfm/sfm/mapper/bundle_adjustment.py:104) and lets Ceres order the parameter blocks and
// analyse the Schur structure itself.  Here that analysis is this file: the two observation orders (counting
// sorts), the block-CSR structure of the reduced camera system, and the Schur work lists, multi-threaded.
// Everything is index work: bit-exact by construction, and checked against the numpy statement of the same
// layout (engine.build_layout_numpy) in tests/test_host_layout.py.
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <stdio.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

#include "../../include/mpba.h"

struct mpba_host_layout {
  std::vector<int32_t> pt_ptr, pm_img, pm_src, img_ptr, cm_pt, cm_src, cm_pm;
  std::vector<int32_t> chunk_img, chunk_begin, chunk_end, img_chunk_ptr;
  std::vector<int32_t> s_row_ptr, s_col, s_diag, s_transpose, s_row, u_block, u_item_ptr, item_a, item_b;
  std::vector<int32_t> track_warp_first;
};

namespace {

int hw_threads() {
  const char* env = getenv("MPBA_HOST_THREADS");
  int n = env ? atoi(env) : (int)std::thread::hardware_concurrency();
  return std::max(1, std::min(n, 64));
}

template <class F>
void parallel_ranges(int64_t n, int threads, F f) {  // f(thread, begin, end) over contiguous ranges
  if (threads <= 1 || n < 4096) {
    f(0, (int64_t)0, n);
    return;
  }
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t) pool.emplace_back([=] { f(t, n * t / threads, n * (t + 1) / threads); });
  for (auto& th : pool) th.join();
}

// stable counting sort of `in` (a permutation of observation indices) by key[in[k]] in [0, nkeys)
void counting_sort(const std::vector<int32_t>& in, const int32_t* key, int32_t nkeys, std::vector<int32_t>& out,
                   std::vector<int64_t>& ptr) {
  const int64_t n = (int64_t)in.size();
  ptr.assign((size_t)nkeys + 1, 0);
  for (int64_t k = 0; k < n; ++k) ++ptr[(size_t)key[in[k]] + 1];
  for (int32_t j = 0; j < nkeys; ++j) ptr[j + 1] += ptr[j];
  std::vector<int64_t> pos(ptr.begin(), ptr.end() - 1);
  out.resize((size_t)n);
  for (int64_t k = 0; k < n; ++k) out[(size_t)pos[key[in[k]]]++] = in[k];
}

// open-addressing set / map of 64-bit keys (keys >= 0); insert() grows the table
struct KeyMap {
  std::vector<int64_t> keys;
  std::vector<int32_t> vals;
  uint64_t mask = 0;
  size_t used = 0;
  void init(size_t expected) {
    size_t cap = 64;
    while (cap < 2 * expected + 16) cap <<= 1;
    keys.assign(cap, -1);
    vals.assign(cap, 0);
    mask = cap - 1;
    used = 0;
  }
  void grow() {
    std::vector<int64_t> ok;
    std::vector<int32_t> ov;
    ok.swap(keys);
    ov.swap(vals);
    init(2 * ok.size());
    for (size_t i = 0; i < ok.size(); ++i)
      if (ok[i] != -1) insert(ok[i], ov[i]);
  }
  static uint64_t mix(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
    return x;
  }
  bool insert(int64_t k, int32_t v) {  // false if present
    uint64_t h = mix((uint64_t)k) & mask;
    while (keys[h] != -1) {
      if (keys[h] == k) return false;
      h = (h + 1) & mask;
    }
    keys[h] = k;
    vals[h] = v;
    if (2 * ++used > keys.size()) grow();
    return true;
  }
  int32_t find(int64_t k) const {
    uint64_t h = mix((uint64_t)k) & mask;
    while (keys[h] != k) h = (h + 1) & mask;
    return vals[h];
  }
};

}  // namespace

extern "C" int32_t mpba_chunk_size(int64_t num_obs) {
  int32_t chunk = MPBA_CHUNK;
  while (chunk > 128 && num_obs / chunk < 592) chunk /= 2;
  return chunk;
}

extern "C" void mpba_host_layout_destroy(mpba_host_layout* L) { delete L; }

extern "C" int mpba_host_layout_create(int32_t C, int32_t P, int32_t N, const int32_t* obs_img, const int32_t* obs_pt,
                                       const uint8_t* pt_const, int32_t structure_only, mpba_host_layout** out,
                                       mpba_host_layout_view* view) {
  if (!out || !view || C < 0 || P < 0 || N < 0 || (N > 0 && (!obs_img || !obs_pt)) || (P > 0 && !pt_const))
    return MPBA_E_INVALID;
  for (int64_t k = 0; k < N; ++k)
    if (obs_img[k] < 0 || obs_img[k] >= C || obs_pt[k] < 0 || obs_pt[k] >= P) return MPBA_E_INVALID;
  // below ~64k observations every phase is shorter than the start of a thread pool
  const int threads = N < 65536 ? 1 : hw_threads();
  mpba_host_layout* L = new mpba_host_layout();
  const bool dbg = getenv("MPBA_DEBUG") != nullptr;
  auto t_prev = std::chrono::steady_clock::now();
  auto lap = [&](const char* what) {
    if (!dbg) return;
    const auto now = std::chrono::steady_clock::now();
    fprintf(stderr, "[mpba] host layout: %-18s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(now - t_prev).count());
    t_prev = now;
  };

  // ---- the two observation orders.  Image order (image, point, natural index): counting sort by image, then every
  // image's segment is sorted on its own (threads take images from a shared counter).  Track order (point, image,
  // natural index): a stable counting sort of the image order by point, threads owning point ranges.
  std::vector<int64_t> ptr_img, ptr_pt;
  {
    std::vector<int32_t> nat((size_t)N), by_img;
    for (int32_t k = 0; k < N; ++k) nat[k] = k;
    counting_sort(nat, obs_img, C, by_img, ptr_img);
    lap("  by image");
    L->cm_src.resize((size_t)N);
    L->cm_pt.resize((size_t)N);
    std::atomic<int32_t> next_img(0);
    // An image's segment of by_img is in natural order (the counting sort is stable), so a STABLE sort by point gives
    // (point, natural index) order: LSD radix sort on the point index, 11 bits per pass (two passes up to 4 M points)
    // instead of a comparison sort of 64-bit keys.
    int pt_bits = 1;
    while (pt_bits < 31 && ((int64_t)1 << pt_bits) < (int64_t)P) ++pt_bits;
    const int passes = (pt_bits + 10) / 11;
    auto sort_images = [&] {
      std::vector<uint64_t> keys, tmp;
      std::vector<int32_t> hist((size_t)2048);
      for (;;) {
        const int32_t i = next_img.fetch_add(1);
        if (i >= C) break;
        const int64_t b = ptr_img[i], e = ptr_img[i + 1];
        const size_t n = (size_t)(e - b);
        keys.resize(n);
        tmp.resize(n);
        for (int64_t k = b; k < e; ++k) keys[k - b] = ((uint64_t)(uint32_t)obs_pt[by_img[k]] << 32) | (uint32_t)by_img[k];
        if (n < 64) {
          std::sort(keys.begin(), keys.end());  // (point, natural index): the same order
        } else {
          for (int pass = 0; pass < passes; ++pass) {
            const int shift = 32 + 11 * pass;
            std::fill(hist.begin(), hist.end(), 0);
            for (size_t k = 0; k < n; ++k) ++hist[(keys[k] >> shift) & 2047u];
            int32_t run = 0;
            for (int d = 0; d < 2048; ++d) {
              const int32_t c = hist[d];
              hist[d] = run;
              run += c;
            }
            for (size_t k = 0; k < n; ++k) tmp[(size_t)hist[(keys[k] >> shift) & 2047u]++] = keys[k];
            keys.swap(tmp);
          }
        }
        for (int64_t k = b; k < e; ++k) {
          L->cm_src[k] = (int32_t)(keys[k - b] & 0xffffffffu);
          L->cm_pt[k] = (int32_t)(keys[k - b] >> 32);
        }
      }
    };
    if (threads > 1 && N >= 4096) {
      std::vector<std::thread> pool;
      for (int t = 0; t < threads; ++t) pool.emplace_back(sort_images);
      for (auto& th : pool) th.join();
    } else {
      sort_images();
    }
  }
  lap("  inside images");
  ptr_pt.assign((size_t)P + 1, 0);
  for (int64_t k = 0; k < N; ++k) ++ptr_pt[(size_t)L->cm_pt[k] + 1];
  for (int32_t p = 0; p < P; ++p) ptr_pt[p + 1] += ptr_pt[p];
  L->pm_src.resize((size_t)N);
  {
    const int T = (N >= 65536) ? threads : 1;
    std::vector<int64_t> pos(ptr_pt.begin(), ptr_pt.end() - 1);
    auto place = [&](int t) {
      const int32_t p0 = (int32_t)((int64_t)P * t / T), p1 = (int32_t)((int64_t)P * (t + 1) / T);
      const int32_t* cm_pt = L->cm_pt.data();
      for (int64_t k = 0; k < N; ++k) {
        const int32_t p = cm_pt[k];
        if (p >= p0 && p < p1) L->pm_src[(size_t)pos[p]++] = L->cm_src[k];
      }
    };
    if (T > 1) {
      std::vector<std::thread> pool;
      for (int t = 0; t < T; ++t) pool.emplace_back(place, t);
      for (auto& th : pool) th.join();
    } else {
      place(0);
    }
  }
  lap("  track order");
  L->pt_ptr.assign(ptr_pt.begin(), ptr_pt.end());
  L->img_ptr.assign(ptr_img.begin(), ptr_img.end());
  L->pm_img.resize((size_t)N);
  L->cm_pm.resize((size_t)N);
  {
    std::vector<int32_t> pm_pos((size_t)N);
    parallel_ranges(N, threads, [&](int, int64_t b, int64_t e) {
      for (int64_t k = b; k < e; ++k) {
        L->pm_img[k] = obs_img[L->pm_src[k]];
        pm_pos[L->pm_src[k]] = (int32_t)k;
      }
    });
    parallel_ranges(N, threads, [&](int, int64_t b, int64_t e) {
      for (int64_t k = b; k < e; ++k) L->cm_pm[k] = pm_pos[L->cm_src[k]];
    });
  }

  lap("observation orders");
  // ---- image-order work chunks
  const int64_t CHUNK = mpba_chunk_size(ptr_img[C]);
  L->img_chunk_ptr.assign((size_t)C + 1, 0);
  for (int32_t i = 0; i < C; ++i) {
    const int64_t cnt = ptr_img[i + 1] - ptr_img[i];
    L->img_chunk_ptr[i + 1] = L->img_chunk_ptr[i] + (int32_t)((cnt + CHUNK - 1) / CHUNK);
  }
  const int32_t nch = L->img_chunk_ptr[C];
  L->chunk_img.resize(nch); L->chunk_begin.resize(nch); L->chunk_end.resize(nch);
  for (int32_t i = 0; i < C; ++i)
    for (int32_t c = L->img_chunk_ptr[i]; c < L->img_chunk_ptr[i + 1]; ++c) {
      const int64_t b = ptr_img[i] + (int64_t)(c - L->img_chunk_ptr[i]) * CHUNK;
      L->chunk_img[c] = i;
      L->chunk_begin[c] = (int32_t)b;
      L->chunk_end[c] = (int32_t)std::min<int64_t>(b + CHUNK, ptr_img[i + 1]);
    }

  // ---- reduced camera system: block-CSR over co-visible image pairs, Schur work lists
  std::vector<int64_t> pair_ptr((size_t)P + 1, 0);
  if (!structure_only)
    for (int32_t p = 0; p < P; ++p) {
      const int64_t ell = pt_const[p] ? 0 : ptr_pt[p + 1] - ptr_pt[p];
      pair_ptr[p + 1] = pair_ptr[p] + ell * (ell + 1) / 2;
    }
  const int64_t n_items = pair_ptr[P];
  if (n_items >= (int64_t)1 << 31) {
    delete L;
    return MPBA_E_INVALID;
  }
  const int32_t* pm_img = L->pm_img.data();
  std::vector<int64_t> all_keys;
  // Up to 2048 images the (image, image) -> block map is a dense C x C table (<= 16 MB, a few hundred KB for the
  // mapper's bundles): marking and look-up are one indexed access, and a row-major scan yields the blocks sorted.
  const bool dense_table = (int64_t)C * C <= ((int64_t)1 << 22);
  std::vector<int32_t> table;
  if (dense_table) {
    table.assign((size_t)C * C, -1);
    std::vector<std::atomic<uint8_t>> mark((size_t)C * C);
    for (auto& m : mark) m.store(0, std::memory_order_relaxed);
    if (n_items > 0)
      parallel_ranges(P, threads, [&](int, int64_t pb, int64_t pe) {
        for (int64_t p = pb; p < pe; ++p) {
          if (pair_ptr[p + 1] == pair_ptr[p]) continue;
          const int64_t s0 = ptr_pt[p], s1 = ptr_pt[p + 1];
          for (int64_t a = s0; a < s1; ++a)
            for (int64_t b = a; b < s1; ++b) {
              const size_t i = (size_t)pm_img[a], j = (size_t)pm_img[b];
              // test first: once a pair is marked its line stays shared between the threads' caches
              if (!mark[i * C + j].load(std::memory_order_relaxed)) {
                mark[i * C + j].store(1, std::memory_order_relaxed);
                mark[j * C + i].store(1, std::memory_order_relaxed);
              }
            }
        }
      });
    for (int64_t i = 0; i < C; ++i) mark[(size_t)(i * C + i)].store(1, std::memory_order_relaxed);
    for (size_t k = 0; k < mark.size(); ++k)
      if (mark[k].load(std::memory_order_relaxed)) {
        table[k] = (int32_t)all_keys.size();
        all_keys.push_back((int64_t)k);
      }
  } else {
    std::vector<std::vector<int64_t>> found((size_t)threads);
    if (n_items > 0)
      parallel_ranges(P, threads, [&](int t, int64_t pb, int64_t pe) {
        KeyMap seen;
        seen.init(1024);
        std::vector<int64_t>& mine = found[t];
        for (int64_t p = pb; p < pe; ++p) {
          if (pair_ptr[p + 1] == pair_ptr[p]) continue;
          const int64_t s0 = ptr_pt[p], s1 = ptr_pt[p + 1];
          for (int64_t a = s0; a < s1; ++a)
            for (int64_t b = a; b < s1; ++b) {
              const int64_t key = (int64_t)pm_img[a] * C + pm_img[b];
              if (seen.insert(key, 0)) mine.push_back(key);
            }
        }
      });
    for (auto& v : found)
      for (int64_t key : v) {
        all_keys.push_back(key);
        const int64_t i = key / C, j = key % C;
        if (i != j) all_keys.push_back(j * C + i);
      }
    for (int64_t i = 0; i < C; ++i) all_keys.push_back(i * ((int64_t)C + 1));
    std::sort(all_keys.begin(), all_keys.end());
    all_keys.erase(std::unique(all_keys.begin(), all_keys.end()), all_keys.end());
  }
  lap("unique blocks");
  const int64_t B = (int64_t)all_keys.size();
  if (B >= (int64_t)1 << 31) {
    delete L;
    return MPBA_E_INVALID;
  }
  KeyMap block_map;
  if (!dense_table) {
    block_map.init((size_t)B);
    for (int64_t b = 0; b < B; ++b) block_map.insert(all_keys[b], (int32_t)b);
  }
  auto block_of_find = [&](int64_t key) -> int32_t { return dense_table ? table[(size_t)key] : block_map.find(key); };
  L->s_row_ptr.assign((size_t)C + 1, 0);
  L->s_col.resize((size_t)B); L->s_row.resize((size_t)B); L->s_transpose.resize((size_t)B); L->s_diag.resize((size_t)C);
  for (int64_t b = 0; b < B; ++b) {
    const int64_t i = all_keys[b] / C, j = all_keys[b] % C;
    L->s_row[b] = (int32_t)i;
    L->s_col[b] = (int32_t)j;
    ++L->s_row_ptr[i + 1];
    L->s_transpose[b] = block_of_find(j * C + i);
    if (i == j) L->s_diag[i] = (int32_t)b;
    if (i <= j) L->u_block.push_back((int32_t)b);
  }
  for (int32_t i = 0; i < C; ++i) L->s_row_ptr[i + 1] += L->s_row_ptr[i];

  // items: generated track by track, placed by a counting sort over the blocks that keeps the track order
  L->item_a.resize((size_t)n_items);
  L->item_b.resize((size_t)n_items);
  std::vector<int64_t> total((size_t)B, 0);
  if (n_items > 0) {
    std::vector<int32_t> item_block((size_t)n_items);
    const int T = (P < 4096) ? 1 : threads;
    std::vector<std::vector<int32_t>> cnt((size_t)T, std::vector<int32_t>((size_t)B, 0));
    auto range = [&](int t, int64_t& pb, int64_t& pe) { pb = (int64_t)P * t / T; pe = (int64_t)P * (t + 1) / T; };
    auto run = [&](auto body) {
      if (T == 1) { body(0); return; }
      std::vector<std::thread> pool;
      for (int t = 0; t < T; ++t) pool.emplace_back([&body, t] { body(t); });
      for (auto& th : pool) th.join();
    };
    run([&](int t) {
      int64_t pb, pe;
      range(t, pb, pe);
      std::vector<int32_t>& c = cnt[t];
      for (int64_t p = pb; p < pe; ++p) {
        if (pair_ptr[p + 1] == pair_ptr[p]) continue;
        const int64_t s0 = ptr_pt[p], s1 = ptr_pt[p + 1];
        int64_t k = pair_ptr[p];
        int64_t last_key = -1;
        int32_t last_blk = 0;
        for (int64_t a = s0; a < s1; ++a)
          for (int64_t b = a; b < s1; ++b, ++k) {
            const int64_t key = (int64_t)pm_img[a] * C + pm_img[b];
            if (key != last_key) { last_key = key; last_blk = block_of_find(key); }
            item_block[k] = last_blk;
            ++c[last_blk];
          }
      }
    });
    lap("item blocks");
    // first slot of (block, thread): blocks in order, threads in order inside a block
    std::vector<std::vector<int64_t>> start((size_t)T, std::vector<int64_t>((size_t)B, 0));
    int64_t run_total = 0;
    for (int64_t b = 0; b < B; ++b)
      for (int t = 0; t < T; ++t) {
        start[t][b] = run_total;
        run_total += cnt[t][b];
        total[b] += cnt[t][b];
      }
    run([&](int t) {
      int64_t pb, pe;
      range(t, pb, pe);
      std::vector<int64_t>& st = start[t];
      for (int64_t p = pb; p < pe; ++p) {
        if (pair_ptr[p + 1] == pair_ptr[p]) continue;
        const int64_t s0 = ptr_pt[p], s1 = ptr_pt[p + 1];
        int64_t k = pair_ptr[p];
        for (int64_t a = s0; a < s1; ++a)
          for (int64_t b = a; b < s1; ++b, ++k) {
            const int64_t dst = st[item_block[k]]++;
            L->item_a[dst] = (int32_t)a;
            L->item_b[dst] = (int32_t)b;
          }
      }
    });
  }
  lap("item placement");
  L->u_item_ptr.assign(L->u_block.size() + 1, 0);
  for (size_t u = 0; u < L->u_block.size(); ++u) L->u_item_ptr[u + 1] = L->u_item_ptr[u] + (int32_t)total[L->u_block[u]];

  // ---- warp groups of whole tracks
  L->track_warp_first.resize((size_t)P + 1);
  int32_t nw = 0;
  if (mpba_pack_tracks(L->pt_ptr.data(), 0, P, L->track_warp_first.data(), &nw) != MPBA_OK) {
    delete L;
    return MPBA_E_INVALID;
  }
  L->track_warp_first.resize((size_t)nw + 1);

  memset(view, 0, sizeof(*view));
  view->num_images = C; view->num_points = P; view->num_obs = N;
  view->num_chunks = nch; view->num_blocks = (int32_t)B; view->num_upper = (int32_t)L->u_block.size();
  view->num_track_warps = nw; view->num_items = n_items;
  view->pt_ptr = L->pt_ptr.data(); view->pm_img = L->pm_img.data(); view->pm_src = L->pm_src.data();
  view->img_ptr = L->img_ptr.data(); view->cm_pt = L->cm_pt.data(); view->cm_src = L->cm_src.data();
  view->cm_pm = L->cm_pm.data();
  view->chunk_img = L->chunk_img.data(); view->chunk_begin = L->chunk_begin.data(); view->chunk_end = L->chunk_end.data();
  view->img_chunk_ptr = L->img_chunk_ptr.data();
  view->s_row_ptr = L->s_row_ptr.data(); view->s_col = L->s_col.data(); view->s_diag = L->s_diag.data();
  view->s_transpose = L->s_transpose.data(); view->s_row = L->s_row.data();
  view->u_block = L->u_block.data(); view->u_item_ptr = L->u_item_ptr.data();
  view->item_a = L->item_a.data(); view->item_b = L->item_b.data();
  view->track_warp_first = L->track_warp_first.data();
  *out = L;
  return MPBA_OK;
}

// out_xy[k] = obs_xy[src[k]], out_depth[r][k] = {d, w, a}[r][src[k]]: the value arrays of one observation order
extern "C" int mpba_host_gather(int32_t N, const int32_t* src, const double* obs_xy, const double* d, const double* w,
                                const double* a, double* out_xy, double* out_depth) {
  if (N < 0 || (N > 0 && (!src || !obs_xy || !d || !w || !a || !out_xy || !out_depth))) return MPBA_E_INVALID;
  parallel_ranges(N, hw_threads(), [&](int, int64_t b, int64_t e) {
    for (int64_t k = b; k < e; ++k) {
      const int64_t j = src[k];
      out_xy[2 * k] = obs_xy[2 * j];
      out_xy[2 * k + 1] = obs_xy[2 * j + 1];
      out_depth[k] = d[j];
      out_depth[(int64_t)N + k] = w[j];
      out_depth[2 * (int64_t)N + k] = a[j];
    }
  });
  return MPBA_OK;
}
