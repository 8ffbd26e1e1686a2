// b200ps N6 — native libsvm batch loader: reader threads -> packed CSR batches in a ring of pinned buffers.
//
// The reference's input side is Python threads that read lines through tf.gfile, keep or drop each
// line with a Bernoulli test, and either collect Sample objects (mode 'all') or push raw text into a
// TF FIFOQueue that tf.train.batch drains (mode 'queue'); every batch is then re-formatted by CPython
// loops before sess.run sees it (lr2.py:87-155,159-169,69-84).  Here the whole path — file read, line
// split, number parsing, sampling, batching — is native and ends in a buffer the GPU can take as is:
//
//   reader thread t  : files[t::threads] (lr2.py:232-233), 1 MiB reads, lines parsed in place
//   batch builder    : each thread fills a free ring buffer; B rows -> published to the ready queue
//   consumer         : next() hands out a ready buffer, release() returns it to the free list
//
// A buffer is one *packed batch*: a 64-byte header, then row_ptr int32[B+1], fid int64[cap_nnz],
// fval f32[cap_nnz], labels f32[B], each section 16-byte aligned at a FIXED offset, so one H2D copy of
// the buffer feeds the sparse kernels and a captured CUDA graph can keep pointing at the same device
// offsets (HostFeeder).  Buffers are pinned (cudaHostAlloc) when a CUDA device is present.
//
// Batches always hold exactly B rows ("Batch size. Must divide evenly into the dataset sizes",
// lr2.py:25): a thread's incomplete batch at the end of its data is dropped and counted.
#include <cuda_runtime.h>

#include <atomic>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../common.h"
#include "fast_parse.h"

using namespace b200ps;

namespace {

constexpr uint32_t kMagic = 0xB200C5A0u;

struct Layout {
  int64_t bytes, off_rowptr, off_fid, off_fval, off_labels, cap_rows, cap_nnz;
};

inline int64_t align16(int64_t v) { return (v + 15) & ~int64_t(15); }

Layout make_layout(int64_t rows, int64_t cap_nnz) {
  Layout l;
  l.cap_rows = rows;
  l.cap_nnz = cap_nnz;
  l.off_rowptr = 64;
  l.off_fid = align16(l.off_rowptr + 4 * (rows + 1));
  l.off_fval = align16(l.off_fid + 8 * cap_nnz);
  l.off_labels = align16(l.off_fval + 4 * cap_nnz);
  l.bytes = align16(l.off_labels + 4 * rows);
  return l;
}

struct Header {          // first 64 bytes of a packed batch
  uint32_t magic;
  int32_t rows, nnz, cap_rows, cap_nnz;
  int32_t off_rowptr, off_fid, off_fval, off_labels, bytes;
  int32_t epoch, thread;
  int64_t seq;
  int32_t pad[2];
};
static_assert(sizeof(Header) == 64, "packed batch header is 64 bytes");

struct Loader {
  std::vector<std::string> files;
  int threads = 1, epochs = 1;
  double sampling_rate = 1.0;
  uint64_t seed = 0;
  Layout lay{};
  bool pinned = false;
  std::vector<char*> buffers;
  std::vector<char> buffer_pinned;
  std::mutex mu;
  std::condition_variable cv_free, cv_ready;
  std::deque<char*> free_q, ready_q;
  int live_threads = 0;
  bool stop = false;
  std::vector<std::thread> workers;
  std::atomic<int64_t> lines{0}, kept{0}, sampled_out{0}, malformed{0}, dropped_nnz{0}, batches{0}, dropped_tail{0};
  std::atomic<int64_t> seq{0};
};

// xorshift64* — one stream per reader thread
inline uint64_t rng_next(uint64_t& s) {
  s ^= s >> 12; s ^= s << 25; s ^= s >> 27;
  return s * 2685821657736338717ull;
}

// Decimal float: the exact fast path of fast_parse.h; anything fancier (exponent, inf, nan, hex, long digit strings)
// goes through strtof.
inline bool parse_float(const char*& q, const char* le, float* out) {
  if (b200ps_parse::fast_float(q, le, out)) return true;
  // strtof skips leading white space (new lines included) and needs a terminator: refuse an empty value
  // here, and a result that ran past the line; every line handed to add_line ends in '\n' or NUL
  if (q >= le || *q == ' ' || *q == '\t' || *q == '\r' || *q == '\n') return false;
  char* e = nullptr;
  const float v = strtof(q, &e);
  if (e == q || e > le) return false;
  *out = v;
  q = e;
  return true;
}

struct Builder {          // a batch under construction in a ring buffer
  char* buf = nullptr;
  int32_t* row_ptr = nullptr;
  int64_t* fid = nullptr;
  float* fval = nullptr;
  float* labels = nullptr;
  int rows = 0;
  int64_t nnz = 0;
};

bool acquire(Loader& L, Builder& b) {
  std::unique_lock<std::mutex> lk(L.mu);
  L.cv_free.wait(lk, [&] { return L.stop || !L.free_q.empty(); });
  if (L.stop) return false;
  b.buf = L.free_q.front();
  L.free_q.pop_front();
  lk.unlock();
  b.row_ptr = reinterpret_cast<int32_t*>(b.buf + L.lay.off_rowptr);
  b.fid = reinterpret_cast<int64_t*>(b.buf + L.lay.off_fid);
  b.fval = reinterpret_cast<float*>(b.buf + L.lay.off_fval);
  b.labels = reinterpret_cast<float*>(b.buf + L.lay.off_labels);
  b.rows = 0;
  b.nnz = 0;
  b.row_ptr[0] = 0;
  return true;
}

void publish(Loader& L, Builder& b, int epoch, int tid) {
  Header* h = reinterpret_cast<Header*>(b.buf);
  h->magic = kMagic; h->rows = b.rows; h->nnz = (int32_t)b.nnz;
  h->cap_rows = (int32_t)L.lay.cap_rows; h->cap_nnz = (int32_t)L.lay.cap_nnz;
  h->off_rowptr = (int32_t)L.lay.off_rowptr; h->off_fid = (int32_t)L.lay.off_fid;
  h->off_fval = (int32_t)L.lay.off_fval; h->off_labels = (int32_t)L.lay.off_labels; h->bytes = (int32_t)L.lay.bytes;
  h->epoch = epoch; h->thread = tid; h->seq = L.seq.fetch_add(1);
  h->pad[0] = h->pad[1] = 0;
  {
    std::lock_guard<std::mutex> lk(L.mu);
    L.ready_q.push_back(b.buf);
  }
  L.cv_ready.notify_one();
  L.batches.fetch_add(1);
  b.buf = nullptr;
}

// One line [p, le) appended to the batch.  Returns false (row not added) for a malformed line or when
// the row would overflow the non-zero capacity of the buffer.
bool add_line(Loader& L, Builder& b, const char* p, const char* le) {
  char* e = nullptr;
  if (p >= le || *p == ' ' || *p == '\t' || *p == '\r') { L.malformed.fetch_add(1); return false; }
  const long lab = strtol(p, &e, 10);
  if (e == p || e > le) { L.malformed.fetch_add(1); return false; }
  const char* q = e;
  int64_t z = b.nnz;
  while (q < le) {
    while (q < le && (*q == ' ' || *q == '\t' || *q == '\r')) ++q;
    if (q >= le) break;
    int64_t idx = 0;
    const char* d0 = q;
    while (q < le && *q >= '0' && *q <= '9') { idx = idx * 10 + (*q - '0'); ++q; }
    if (q == d0 || q >= le || *q != ':') { L.malformed.fetch_add(1); return false; }
    ++q;
    float val;
    if (!parse_float(q, le, &val)) { L.malformed.fetch_add(1); return false; }
    if (z >= L.lay.cap_nnz) { L.dropped_nnz.fetch_add(1); return false; }
    b.fid[z] = idx;
    b.fval[z] = val;
    ++z;
  }
  b.labels[b.rows] = (float)lab;
  b.nnz = z;
  ++b.rows;
  b.row_ptr[b.rows] = (int32_t)z;
  return true;
}

void reader(Loader* Lp, int tid) {
  Loader& L = *Lp;
  uint64_t rs = (L.seed + 0x9E3779B97F4A7C15ull * (uint64_t)(tid + 1)) | 1ull;
  const uint64_t keep_below = L.sampling_rate >= 1.0 ? ~0ull : (uint64_t)(L.sampling_rate * 18446744073709551615.0);
  Builder b;
  std::vector<char> chunk(1 << 20);
  std::string carry;
  bool ok = true;
  for (int epoch = 0; ok && (L.epochs <= 0 || epoch < L.epochs); ++epoch) {
    for (size_t fi = tid; ok && fi < L.files.size(); fi += L.threads) {
      FILE* f = fopen(L.files[fi].c_str(), "rb");
      if (!f) { L.malformed.fetch_add(1); continue; }
      carry.clear();
      for (;;) {
        const size_t n = fread(chunk.data(), 1, chunk.size(), f);
        const bool eof = n == 0;
        const char* p = chunk.data();
        const char* end = p + n;
        while (ok) {
          const char* nl = p < end ? static_cast<const char*>(memchr(p, '\n', end - p)) : nullptr;
          const char *ls, *le;
          if (nl) {
            if (!carry.empty()) { carry.append(p, nl - p); ls = carry.data(); le = ls + carry.size(); }
            else { ls = p; le = nl; }
          } else if (eof && !carry.empty()) {
            ls = carry.data(); le = ls + carry.size();
          } else {
            break;
          }
          if (le - ls >= 2) {                                   // lr2.py:105 skips shorter lines
            L.lines.fetch_add(1, std::memory_order_relaxed);
            if (L.sampling_rate >= 1.0 || rng_next(rs) <= keep_below) {
              if (!b.buf && !acquire(L, b)) { ok = false; break; }
              if (add_line(L, b, ls, le)) L.kept.fetch_add(1, std::memory_order_relaxed);
              if (b.rows == (int)L.lay.cap_rows) publish(L, b, epoch, tid);
            } else {
              L.sampled_out.fetch_add(1, std::memory_order_relaxed);
            }
          }
          if (nl) { carry.clear(); p = nl + 1; }
          else { carry.clear(); break; }
        }
        if (!ok || eof) break;
        if (p < end) carry.append(p, end - p);                  // partial last line of the chunk
      }
      fclose(f);
    }
  }
  if (b.buf) {                                                  // incomplete batch: dropped
    L.dropped_tail.fetch_add(b.rows);
    std::lock_guard<std::mutex> lk(L.mu);
    L.free_q.push_back(b.buf);
  }
  {
    std::lock_guard<std::mutex> lk(L.mu);
    --L.live_threads;
  }
  L.cv_ready.notify_all();
}

}  // namespace

extern "C" int b200ps_loader_create(void** out, const char* const* files, int32_t nfiles, int32_t threads,
                                    int32_t batch_rows, int64_t cap_nnz, int32_t ring, double sampling_rate,
                                    uint64_t seed, int32_t epochs, int32_t pinned) {
  B200PS_REQUIRE(out && files && nfiles >= 0 && threads >= 1 && batch_rows >= 1 && cap_nnz >= 1 && ring >= 2,
                 "loader_create: bad arguments");
  B200PS_REQUIRE(sampling_rate > 0.0 && sampling_rate <= 1.0, "loader_create: sampling_rate must be in (0, 1]");
  Loader* L = new Loader();
  for (int i = 0; i < nfiles; ++i) L->files.emplace_back(files[i]);
  L->threads = threads; L->epochs = epochs; L->sampling_rate = sampling_rate; L->seed = seed;
  L->lay = make_layout(batch_rows, cap_nnz);
  if (L->lay.bytes >= (int64_t(1) << 31)) {
    delete L;
    return set_error(B200PS_ERR_INVALID, "loader_create: packed batch larger than 2 GiB");
  }
  L->pinned = pinned != 0;
  for (int i = 0; i < ring; ++i) {
    char* p = nullptr;
    bool is_pinned = false;
    if (L->pinned) {
      if (cudaHostAlloc(reinterpret_cast<void**>(&p), L->lay.bytes, cudaHostAllocPortable) == cudaSuccess) {
        is_pinned = true;
      } else {
        cudaGetLastError();          // no device / no driver: plain memory (CPU mode and the CPU tests)
        L->pinned = false;
        p = nullptr;
      }
    }
    if (!p) p = static_cast<char*>(aligned_alloc(256, (size_t)((L->lay.bytes + 255) & ~int64_t(255))));
    if (!p) {
      for (size_t k = 0; k < L->buffers.size(); ++k) {
        if (L->buffer_pinned[k]) cudaFreeHost(L->buffers[k]); else free(L->buffers[k]);
      }
      delete L;
      return set_error(B200PS_ERR_INVALID, "loader_create: out of memory");
    }
    memset(p, 0, L->lay.bytes);
    L->buffers.push_back(p);
    L->buffer_pinned.push_back(is_pinned ? 1 : 0);
    L->free_q.push_back(p);
  }
  L->live_threads = threads;
  for (int t = 0; t < threads; ++t) L->workers.emplace_back(reader, L, t);
  *out = L;
  return 0;
}

// out[7] = bytes, off_rowptr, off_fid, off_fval, off_labels, cap_rows, cap_nnz; *pinned = buffers are page-locked
extern "C" int b200ps_loader_layout(void* loader, int64_t* out, int32_t* pinned) {
  Loader* L = static_cast<Loader*>(loader);
  B200PS_REQUIRE(L && out, "loader_layout: null argument");
  out[0] = L->lay.bytes; out[1] = L->lay.off_rowptr; out[2] = L->lay.off_fid; out[3] = L->lay.off_fval;
  out[4] = L->lay.off_labels; out[5] = L->lay.cap_rows; out[6] = L->lay.cap_nnz;
  if (pinned) *pinned = L->pinned ? 1 : 0;
  return 0;
}

// Layout of a packed batch without a loader (for callers that pack their own batches).
extern "C" int b200ps_packed_layout(int64_t batch_rows, int64_t cap_nnz, int64_t* out) {
  B200PS_REQUIRE(out && batch_rows >= 1 && cap_nnz >= 1, "packed_layout: bad arguments");
  const Layout l = make_layout(batch_rows, cap_nnz);
  out[0] = l.bytes; out[1] = l.off_rowptr; out[2] = l.off_fid; out[3] = l.off_fval; out[4] = l.off_labels;
  out[5] = l.cap_rows; out[6] = l.cap_nnz;
  return 0;
}

// 0: *buf is the next ready batch (give it back with release); 1: end of data; 2: timeout.
extern "C" int b200ps_loader_next(void* loader, void** buf, int32_t timeout_ms) {
  Loader* L = static_cast<Loader*>(loader);
  B200PS_REQUIRE(L && buf, "loader_next: null argument");
  std::unique_lock<std::mutex> lk(L->mu);
  auto ready = [&] { return !L->ready_q.empty() || L->live_threads == 0 || L->stop; };
  if (timeout_ms < 0) L->cv_ready.wait(lk, ready);
  else if (!L->cv_ready.wait_for(lk, std::chrono::milliseconds(timeout_ms), ready)) return 2;
  if (L->ready_q.empty()) return 1;
  *buf = L->ready_q.front();
  L->ready_q.pop_front();
  return 0;
}

extern "C" int b200ps_loader_release(void* loader, void* buf) {
  Loader* L = static_cast<Loader*>(loader);
  B200PS_REQUIRE(L && buf, "loader_release: null argument");
  {
    std::lock_guard<std::mutex> lk(L->mu);
    L->free_q.push_back(static_cast<char*>(buf));
  }
  L->cv_free.notify_one();
  return 0;
}

// out[7] = lines seen, rows kept, rows sampled out, malformed lines, rows dropped for the nnz cap,
//          batches published, rows dropped in incomplete final batches
extern "C" int b200ps_loader_stats(void* loader, int64_t* out) {
  Loader* L = static_cast<Loader*>(loader);
  B200PS_REQUIRE(L && out, "loader_stats: null argument");
  out[0] = L->lines; out[1] = L->kept; out[2] = L->sampled_out; out[3] = L->malformed; out[4] = L->dropped_nnz;
  out[5] = L->batches; out[6] = L->dropped_tail;
  return 0;
}

extern "C" int b200ps_loader_destroy(void* loader) {
  Loader* L = static_cast<Loader*>(loader);
  if (!L) return 0;
  {
    std::lock_guard<std::mutex> lk(L->mu);
    L->stop = true;
  }
  L->cv_free.notify_all();
  L->cv_ready.notify_all();
  for (auto& w : L->workers) w.join();
  for (size_t i = 0; i < L->buffers.size(); ++i) {
    if (L->buffer_pinned[i]) cudaFreeHost(L->buffers[i]); else free(L->buffers[i]);
  }
  delete L;
  return 0;
}
